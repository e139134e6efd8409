#pragma once
/* Dim3: 3-component int64 vector used for sizes, positions and direction vectors.

   Capability parity with the reference's include/stencil/dim3.hpp:25-314 (arithmetic, wrap,
   lexicographic ordering, make_block_dim, stream output), written fresh for this library.
   Deliberate differences: max()/operator!=/const operator[] are correct here (the reference has
   typos in all three). */

#include <algorithm>
#include <cassert>
#include <cstdint>
#include <ostream>

#if defined(__CUDACC__)
#define STENCIL_HD __host__ __device__
#define STENCIL_INLINE __forceinline__
#else
#define STENCIL_HD
#define STENCIL_INLINE inline
#endif

/* smallest power of two >= x (x >= 1) */
inline int64_t nextPowerOfTwo(int64_t x) {
  int64_t p = 1;
  while (p < x) p <<= 1;
  return p;
}

class Dim3 {
public:
  int64_t x, y, z;

  STENCIL_HD constexpr Dim3() : x(0), y(0), z(0) {}
  STENCIL_HD constexpr Dim3(int64_t _x, int64_t _y, int64_t _z) : x(_x), y(_y), z(_z) {}
#if defined(__CUDACC__)
  STENCIL_HD Dim3(const dim3 &d) : x(d.x), y(d.y), z(d.z) {}
  STENCIL_HD operator dim3() const { return dim3((unsigned)x, (unsigned)y, (unsigned)z); }
#endif

  STENCIL_HD int64_t &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
  STENCIL_HD const int64_t &operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }

  /* number of points in a box of this extent */
  STENCIL_HD int64_t flatten() const { return x * y * z; }

  STENCIL_HD int64_t max() const { return x > y ? (x > z ? x : z) : (y > z ? y : z); }
  STENCIL_HD int64_t min() const { return x < y ? (x < z ? x : z) : (y < z ? y : z); }

  /* x-major lexicographic order: this ordering defines message order inside packed buffers */
  STENCIL_HD bool operator<(const Dim3 &r) const {
    if (x != r.x) return x < r.x;
    if (y != r.y) return y < r.y;
    return z < r.z;
  }
  STENCIL_HD bool operator==(const Dim3 &r) const { return x == r.x && y == r.y && z == r.z; }
  STENCIL_HD bool operator!=(const Dim3 &r) const { return !(*this == r); }

  STENCIL_HD bool all_lt(int64_t v) const { return x < v && y < v && z < v; }
  STENCIL_HD bool all_gt(int64_t v) const { return x > v && y > v && z > v; }
  STENCIL_HD bool all_ge(int64_t v) const { return x >= v && y >= v && z >= v; }
  STENCIL_HD bool any_lt(int64_t v) const { return x < v || y < v || z < v; }
  STENCIL_HD bool any_gt(int64_t v) const { return x > v || y > v || z > v; }
  STENCIL_HD bool all_lt(const Dim3 &r) const { return x < r.x && y < r.y && z < r.z; }
  STENCIL_HD bool all_ge(const Dim3 &r) const { return x >= r.x && y >= r.y && z >= r.z; }

#define STENCIL_DIM3_OP(op)                                                                                            \
  STENCIL_HD Dim3 &operator op##=(const Dim3 &r) {                                                                     \
    x op## = r.x;                                                                                                      \
    y op## = r.y;                                                                                                      \
    z op## = r.z;                                                                                                      \
    return *this;                                                                                                      \
  }                                                                                                                    \
  STENCIL_HD Dim3 operator op(const Dim3 &r) const {                                                                   \
    Dim3 t(*this);                                                                                                     \
    t op## = r;                                                                                                        \
    return t;                                                                                                          \
  }                                                                                                                    \
  STENCIL_HD Dim3 &operator op##=(int64_t s) {                                                                         \
    x op## = s;                                                                                                        \
    y op## = s;                                                                                                        \
    z op## = s;                                                                                                        \
    return *this;                                                                                                      \
  }                                                                                                                    \
  STENCIL_HD Dim3 operator op(int64_t s) const {                                                                       \
    Dim3 t(*this);                                                                                                     \
    t op## = s;                                                                                                        \
    return t;                                                                                                          \
  }
  STENCIL_DIM3_OP(+)
  STENCIL_DIM3_OP(-)
  STENCIL_DIM3_OP(*)
  STENCIL_DIM3_OP(/)
  STENCIL_DIM3_OP(%)
#undef STENCIL_DIM3_OP

  STENCIL_HD Dim3 operator-() const { return Dim3(-x, -y, -z); }

  /* periodic wrap into [0, lims); mutates and returns *this (reference semantics) */
  STENCIL_HD Dim3 wrap(const Dim3 &lims) {
    x = ((x % lims.x) + lims.x) % lims.x;
    y = ((y % lims.y) + lims.y) % lims.y;
    z = ((z % lims.z) + lims.z) % lims.z;
    return *this;
  }
  STENCIL_HD Dim3 wrapped(const Dim3 &lims) const {
    Dim3 t(*this);
    return t.wrap(lims);
  }

  /* Shape up to `threads` (<=1024) threads like `extent`: fill x first, then y, then z, each with the
     next power of two of the extent; z is capped at 64 (CUDA block limit). */
  static Dim3 make_block_dim(const Dim3 &extent, int64_t threads) {
    threads = std::min<int64_t>(threads, 1024);
    Dim3 b;
    b.x = std::max<int64_t>(1, std::min(threads, nextPowerOfTwo(extent.x)));
    threads /= b.x;
    b.y = std::max<int64_t>(1, std::min(threads, nextPowerOfTwo(extent.y)));
    threads /= b.y;
    b.z = std::max<int64_t>(1, std::min(threads, nextPowerOfTwo(extent.z)));
    b.z = std::min<int64_t>(b.z, 64);
    return b;
  }
};

inline std::ostream &operator<<(std::ostream &os, const Dim3 &d) {
  return os << '[' << d.x << ',' << d.y << ',' << d.z << ']';
}
