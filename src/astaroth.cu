/* AstarothSim engine and kernels (N4/N5 radius-3 multi-field stencil, N9 init).
   Replaces reference kernels K10/K11 (bin/astaroth_sim.cu:14-83). See include/stencil/astaroth.hpp.

   Kernel shape: a CTA is 32 lanes x 8 rows, a lane owns one 16-byte vector of cells (4 fp32 / 2 fp64) and marches
   along z with the seven centre values z-3..z+3 rotating through registers, so the z direction costs one load per
   plane; x and y neighbours are 16-byte loads that hit L1 (rows are 128-byte aligned). All fields go in one launch
   (blockIdx.z = field x z-chunk). That is the plain kernel (thin regions, Average7); the Star19 hot path is the
   TMA-pipelined kernel further down. Schedules: the reference's interior / exchange() / exterior with a host
   synchronisation per iteration, or (default with the TMA kernel) exchange_async() + one whole-region launch per
   subdomain chained by events, with no host synchronisation in the loop. Folding the push into the kernel as the
   Jacobi engine does is the next step. */
#include "stencil/nvtx.hpp"
#include "stencil/astaroth.hpp"

#include <cmath>
#include <cstring>
#include <limits>
#include <set>

#include "stencil/copy_plan.hpp"
#include "stencil/detail/launch_pool.hpp"
#include "stencil/detail/tma.cuh"

namespace stencil {

namespace {

constexpr int kMaxFields = 16, kLanes = 32, kRows = 8, kR = 3;

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int d) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != d) CUDA_RUNTIME(cudaSetDevice(d));
  }
  ~DevGuard() { cudaSetDevice(prev); }
};

template <typename T> struct Vec;
template <> struct Vec<float> {
  typedef float4 type;
  static constexpr int n = 4;
};
template <> struct Vec<double> {
  typedef double2 type;
  static constexpr int n = 2;
};

template <typename T> struct FieldParams {
  const T *src[kMaxFields]; // compute cell (0,0,0) of curr, per field
  T *dst[kMaxFields];
  long long pitchX, pitchXY;
  int lox, loy, loz, hix, hiy, hiz; // compute coordinates, half open
  int tileZ, nChunks;
  T dt;
};

/* 6th-order second-derivative weights: c0 = -49/18 per axis, then 3/2, -3/20, 1/90 */
template <typename T> __host__ __device__ inline T star19(const T c, const T xm[3], const T xp[3], const T ym[3], const T yp[3], const T zm[3], const T zp[3], T dt) {
  const T c1 = T(1.5), c2 = T(-0.15), c3 = T(1.0 / 90.0), c0 = T(-49.0 / 18.0);
  T lap = T(3) * c0 * c;
  lap += c1 * (((((xm[0] + xp[0]) + ym[0]) + yp[0]) + zm[0]) + zp[0]);
  lap += c2 * (((((xm[1] + xp[1]) + ym[1]) + yp[1]) + zm[1]) + zp[1]);
  lap += c3 * (((((xm[2] + xp[2]) + ym[2]) + yp[2]) + zm[2]) + zp[2]);
  return c + dt * lap;
}

/* the reference's stand-in: (-x + -y + -z + +x + +y + +z) / 6 in that order */
template <typename T> __host__ __device__ inline T average7(T xm, T ym, T zm, T xp, T yp, T zp) {
  T v = 0;
  v += xm;
  v += ym;
  v += zm;
  v += xp;
  v += yp;
  v += zp;
  return v / T(6);
}

template <typename T, bool STAR> __global__ void __launch_bounds__(kLanes *kRows) astaroth_kernel(const __grid_constant__ FieldParams<T> P) {
  typedef typename Vec<T>::type V;
  constexpr int N = Vec<T>::n;
  const int field = blockIdx.z / P.nChunks, chunk = blockIdx.z % P.nChunks;
  const int x0 = (P.lox / N) * N + (blockIdx.x * kLanes + threadIdx.x) * N;
  const int y = P.loy + blockIdx.y * kRows + threadIdx.y;
  const int z0 = P.loz + chunk * P.tileZ, z1 = min(z0 + P.tileZ, P.hiz);
  if (y >= P.hiy || x0 >= P.hix || x0 + N <= P.lox || z0 >= z1) return;
  const T *s = P.src[field] + (long long)y * P.pitchX + x0;
  T *d = P.dst[field] + (long long)y * P.pitchX + x0;
  auto ldv = [](const T *p, T out[N]) {
    const V v = *reinterpret_cast<const V *>(p);
    std::memcpy(out, &v, sizeof(V));
  };
  /* centre column z-3 .. z+3 (STAR) or z-1 .. z+1 (average) */
  T q[7][N];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    if (STAR || (k >= 2 && k <= 4)) ldv(s + (long long)(z0 + k - 3) * P.pitchXY, q[k]);
  }
  for (int z = z0; z < z1; ++z) {
    const T *sp = s + (long long)z * P.pitchXY;
    T out[N];
    if (STAR) {
      /* x neighbours: the vectors left and right of mine (and one further for fp64) */
      T xl[2 * N], xr[2 * N];
      ldv(sp - N, xl + N);
      ldv(sp + N, xr);
      if (N < 3) {
        ldv(sp - 2 * N, xl);
        ldv(sp + 2 * N, xr + N);
      }
      T ym[3][N], yp[3][N];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        ldv(sp - (long long)(k + 1) * P.pitchX, ym[k]);
        ldv(sp + (long long)(k + 1) * P.pitchX, yp[k]);
      }
#pragma unroll
      for (int i = 0; i < N; ++i) {
        /* cells in x order around cell i: window = [xl | q[3] | xr] */
        T win[4 * N + N];
#pragma unroll
        for (int j = 0; j < 2 * N; ++j) win[j] = xl[j];
#pragma unroll
        for (int j = 0; j < N; ++j) win[2 * N + j] = q[3][j];
#pragma unroll
        for (int j = 0; j < 2 * N; ++j) win[3 * N + j] = xr[j];
        const int c = 2 * N + i;
        const T xm[3] = {win[c - 1], win[c - 2], win[c - 3]}, xp[3] = {win[c + 1], win[c + 2], win[c + 3]};
        const T ymv[3] = {ym[0][i], ym[1][i], ym[2][i]}, ypv[3] = {yp[0][i], yp[1][i], yp[2][i]};
        const T zm[3] = {q[2][i], q[1][i], q[0][i]}, zp[3] = {q[4][i], q[5][i], q[6][i]};
        out[i] = star19<T>(q[3][i], xm, xp, ymv, ypv, zm, zp, P.dt);
      }
    } else {
      T l[N], r[N], ym[N], yp[N];
      ldv(sp - N, l);
      ldv(sp + N, r);
      ldv(sp - P.pitchX, ym);
      ldv(sp + P.pitchX, yp);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const T xm = i == 0 ? l[N - 1] : q[3][i - 1], xp = i == N - 1 ? r[0] : q[3][i + 1];
        out[i] = average7<T>(xm, ym[i], q[2][i], xp, yp[i], q[4][i]);
      }
    }
    T *dp = d + (long long)z * P.pitchXY;
    if (x0 >= P.lox && x0 + N <= P.hix) {
      V v;
      std::memcpy(&v, out, sizeof(V));
      *reinterpret_cast<V *>(dp) = v;
    } else {
#pragma unroll
      for (int i = 0; i < N; ++i)
        if (x0 + i >= P.lox && x0 + i < P.hix) dp[i] = out[i];
    }
    /* rotate the centre column and fetch the next plane */
#pragma unroll
    for (int k = (STAR ? 0 : 2); k < (STAR ? 6 : 4); ++k)
#pragma unroll
      for (int i = 0; i < N; ++i) q[k][i] = q[k + 1][i];
    if (z + 1 < z1) ldv(s + (long long)(z + 1 + (STAR ? 3 : 1)) * P.pitchXY, q[STAR ? 6 : 4]);
  }
}

/* Cross-derivative terms (Stencil::Star19Cross). Real Astaroth-style MHD right-hand sides contain mixed second
   derivatives, which read the DIAGONAL neighbours and therefore the edge halos a 19-point star never touches:
     next = curr + dt * ( L6(curr) + eta * ( Dxy + Dxz + Dyz + Dxyz ) )
   Dab = sum_k m_k [ f(+k,+k) - f(+k,-k) - f(-k,+k) + f(-k,-k) ] along the diagonals of the ab plane, k = 1..3, with the
   6th-order weights m = (270, -27, 2) / 720; Dxyz = m_1 * sum of sign(dx dy dz) f(+-1,+-1,+-1) reads the eight corner
   directions. All 26 halo directions take part, so a wrong edge or corner block fails the comparison with the host. */
template <typename T> struct CrossW {
  static constexpr T m1() { return T(270.0 / 720.0); }
  static constexpr T m2() { return T(-27.0 / 720.0); }
  static constexpr T m3() { return T(2.0 / 720.0); }
};

template <typename T, typename At> __host__ __device__ inline T cross_terms(At at) { // at(dx,dy,dz)
  const T m[3] = {CrossW<T>::m1(), CrossW<T>::m2(), CrossW<T>::m3()};
  T s = T(0);
  for (int k = 1; k <= 3; ++k) {
    s += m[k - 1] * (((at(k, k, 0) - at(k, -k, 0)) - at(-k, k, 0)) + at(-k, -k, 0));
    s += m[k - 1] * (((at(k, 0, k) - at(-k, 0, k)) - at(k, 0, -k)) + at(-k, 0, -k));
    s += m[k - 1] * (((at(0, k, k) - at(0, -k, k)) - at(0, k, -k)) + at(0, -k, -k));
  }
  s += m[0] * ((((at(1, 1, 1) - at(1, -1, 1)) - at(-1, 1, 1)) + at(-1, -1, 1)) - (((at(1, 1, -1) - at(1, -1, -1)) - at(-1, 1, -1)) + at(-1, -1, -1)));
  return s;
}

/* Star19Cross without the TMA pipeline: one thread per cell, every point straight from global memory (L1/L2 absorb
   the reuse). Slow and general: thin regions of the overlap schedule, Options::useTma = false, and an independent
   device-side check of the pipelined kernel. */
template <typename T> __global__ void __launch_bounds__(kLanes *kRows) astaroth_cross_kernel(const __grid_constant__ FieldParams<T> P, T eta) {
  const int field = blockIdx.z / P.nChunks, chunk = blockIdx.z % P.nChunks;
  const int x = P.lox + blockIdx.x * kLanes + threadIdx.x;
  const int y = P.loy + blockIdx.y * kRows + threadIdx.y;
  const int z0 = P.loz + chunk * P.tileZ, z1 = min(z0 + P.tileZ, P.hiz);
  if (x >= P.hix || y >= P.hiy) return;
  for (int z = z0; z < z1; ++z) {
    const T *c = P.src[field] + (long long)z * P.pitchXY + (long long)y * P.pitchX + x;
    auto at = [&](int dx, int dy, int dz) { return c[(long long)dz * P.pitchXY + (long long)dy * P.pitchX + dx]; };
    T xm[3], xp[3], ym[3], yp[3], zm[3], zp[3];
    for (int k = 0; k < 3; ++k) {
      xm[k] = at(-k - 1, 0, 0), xp[k] = at(k + 1, 0, 0);
      ym[k] = at(0, -k - 1, 0), yp[k] = at(0, k + 1, 0);
      zm[k] = at(0, 0, -k - 1), zp[k] = at(0, 0, k + 1);
    }
    const T v = star19<T>(c[0], xm, xp, ym, yp, zm, zp, P.dt) + P.dt * (eta * cross_terms<T>(at));
    P.dst[field][(long long)z * P.pitchXY + (long long)y * P.pitchX + x] = v;
  }
}

/* ---------------- TMA-pipelined radius-3 kernel ----------------
   Same structure as the Jacobi kernel (src/jacobi.cu): thread 0 streams the input planes of ONE field of one tile
   into a 6-stage shared-memory ring with cp.async.bulk.tensor.3d, eight warps consume, a lane owns one 16-byte vector
   of cells (N = 4 fp32 / 2 fp64) in each of two rows. Two tile shapes, picked per tile (uniform per CTA):
     wide   32 lanes per row: 32N cells x 16 rows;   narrow   8 lanes per row: 8N cells x 64 rows
   — narrow tiles cover what is left of a row after the last full wide tile (323- and 406-cell rows of the weak-scaling
   rule would otherwise pay a whole tile for 3 / 22 cells). The z direction never goes back to memory: each arriving
   plane adds its "above" contribution to the three pending outputs and its centre stays in registers for the three
   "below" contributions that follow, so a plane is read from shared memory once. */
constexpr int kTRows = 16, kTMargin = 4, kTStages = 6, kTLag = 2, kTThreads = 256;

template <typename T, int LW> struct TmaGeom {
  static constexpr int N = Vec<T>::n;
  static constexpr int kSub = 32 / LW;           // row pairs per warp
  static constexpr int tileX = LW * N;
  static constexpr int tileY = kTRows * kSub;
  static constexpr int boxX = tileX + 2 * kTMargin;
  static constexpr int boxY = tileY + 2 * kR;
  static constexpr int planeBytes = boxX * boxY * int(sizeof(T));
  static constexpr int stageBytes = (planeBytes + 127) / 128 * 128;
  static constexpr int barOffset = kTStages * stageBytes;
  static constexpr int smemBytes = barOffset + 2 * kTStages * 8;
};
template <typename T> constexpr int tma_smem_bytes() {
  return TmaGeom<T, 8>::smemBytes > TmaGeom<T, 32>::smemBytes ? TmaGeom<T, 8>::smemBytes : TmaGeom<T, 32>::smemBytes;
}

/* Fused halo pushes. The kernel keeps all 26 halo directions of `next` current by itself, so in
   steady state no exchange() is left:
     +-x faces       the subdomain is its own x neighbour (precondition of the fused path): the lanes that own the first /
                     last three cells of a row store them a second time, sx cells further / earlier in the same row;
     y, z faces      a tile on such a face copies, after its plane loop, the rows it has just written (still in L2) into
                     the neighbour's halo — peer pointers over NVLink, or plain stores when the neighbour shares the GPU;
     yz edges        tiles on a y AND a z face copy their 3 x 3 row bundle to the diagonal neighbour as well;
     xy, xz edges, corners   every row copied above also carries its own x wrap: cells 0..2 land once more sx cells
                     further, cells sx-3..sx-1 sx cells earlier, in the destination's x halo.
   That is 8 destinations ("push directions") per subdomain. Kernels of neighbouring GPUs hand-shake through one epoch
   flag and one completion counter PER DIRECTION, exactly as the fused Jacobi kernel does per face (src/jacobi.cu):
   a tile waits for the flags of the directions it touches before reading halos, counts itself done on them after a
   GPU-scope fence, and whoever completes a direction raises that neighbour's flag behind one system-scope fence. */
constexpr int kDirs = 8, kPushRows = 4;
enum PushDir { kYM = 0, kYP = 1, kZM = 2, kZP = 3, kYMZM = 4, kYMZP = 5, kYPZM = 6, kYPZP = 7 };
constexpr int kFusedFlagBase = 64; // word offset of the fused-iteration flags inside a control block (as in jacobi.cu)

template <typename T> struct PushParams {
  T *nb[kDirs][kMaxFields];      // my compute cell (x,y,z) lands at nb[d][field] + z*nbPitchXY[d] + y*nbPitchX[d] + x
  long long nbPitchX[kDirs], nbPitchXY[kDirs];
  unsigned *nbFlag[kDirs];       // neighbour's flag to raise (nullptr: ordered by streams on one GPU)
  const unsigned *myFlag[kDirs]; // my flag the neighbour raises
  unsigned dirMask, flagMask;
  unsigned nDir[kDirs]; // CTAs that take part in the hand-shake, per direction
  unsigned *done, *err; // done: kDirs counters
  unsigned long long timeoutNs;
  int sx, sy, sz;
};

__device__ __forceinline__ void st_flag_sys(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

/* Boundary tiles only, after their plane loop: copy the rows of this tile that belong to each push direction from my
   `next` buffer (written a moment ago by this CTA) to the matching halo rows of that neighbour's `next` buffer, x wrap
   included. One warp per row, 16-byte vectors when both sides are aligned. Loads bypass L1 (ld.global.cg). */
template <typename T>
__device__ __noinline__ void push_tile_dirs(const PushParams<T> &Q, int field, unsigned dirs, int xlo, int tw, int ylo, int tr, int z0, int z1,
                                            const T *mine, long long pitchX, long long pitchXY) {
  typedef typename Vec<T>::type V;
  constexpr int N = Vec<T>::n;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = kTThreads / 32;
  const int xhi = min(xlo + tw, Q.sx), yhi = min(ylo + tr, Q.sy);
  for (int d = 0; d < kDirs; ++d) {
    if (!((dirs >> d) & 1u)) continue;
    const int dy = d == kYM || d == kYMZM || d == kYMZP ? -1 : (d == kYP || d == kYPZM || d == kYPZP ? 1 : 0);
    const int dz = d == kZM || d == kYMZM || d == kYPZM ? -1 : (d == kZP || d == kYMZP || d == kYPZP ? 1 : 0);
    int ya = ylo, yb = yhi, za = z0, zb = z1;
    if (dy < 0) yb = min(yb, kR);
    if (dy > 0) ya = max(ya, Q.sy - kR);
    if (dz < 0) zb = min(zb, kR);
    if (dz > 0) za = max(za, Q.sz - kR);
    const int ny = yb - ya, rows = ny * (zb - za);
    if (ny <= 0 || zb <= za) continue;
    T *nb = Q.nb[d][field];
    const long long npx = Q.nbPitchX[d], npxy = Q.nbPitchXY[d];
    /* kPushRows rows per warp and pass, every load of all of them issued before the first store: the loop is bound by
       the L2 round trip of a row, not by bytes */
    const int xe = lane < kR ? lane : Q.sx - 2 * kR + lane;          // x-wrap cell this lane may carry (lanes 0..5)
    const bool ext = lane < 2 * kR && xe >= xlo && xe < xhi;         // ... if this tile owns it
    const int xeDst = lane < kR ? xe + Q.sx : xe - Q.sx;
    for (int r = warp; r < rows; r += kPushRows * nwarps) {
      const T *src[kPushRows];
      T *dst[kPushRows];
      bool on[kPushRows], vec[kPushRows];
      V v[kPushRows];
      T e[kPushRows], tail[kPushRows][N];
      const int x = xlo + lane * N;
#pragma unroll
      for (int k = 0; k < kPushRows; ++k) {
        const int rr = r + k * nwarps;
        on[k] = rr < rows;
        const int z = za + (on[k] ? rr : r) / ny, y = ya + (on[k] ? rr : r) % ny;
        src[k] = mine + (long long)z * pitchXY + (long long)y * pitchX;
        dst[k] = nb + (long long)z * npxy + (long long)y * npx;
        vec[k] = ((reinterpret_cast<unsigned long long>(dst[k] + xlo) | reinterpret_cast<unsigned long long>(src[k] + xlo)) & (sizeof(V) - 1)) == 0;
        v[k] = V();
        e[k] = T(0);
#pragma unroll
        for (int i = 0; i < N; ++i) tail[k][i] = T(0);
        if (on[k]) {
          /* a tile is at most 32 vectors wide: one vector (or its in-range part) per lane */
          if (vec[k] && x + N <= xhi) {
            v[k] = __ldcg(reinterpret_cast<const V *>(src[k] + x));
          } else {
#pragma unroll
            for (int i = 0; i < N; ++i)
              if (x + i < xhi) tail[k][i] = __ldcg(src[k] + x + i);
          }
          if (ext) e[k] = __ldcg(src[k] + xe);
        }
      }
#pragma unroll
      for (int k = 0; k < kPushRows; ++k) {
        if (!on[k]) continue;
        if (vec[k] && x + N <= xhi) {
          *reinterpret_cast<V *>(dst[k] + x) = v[k];
        } else {
#pragma unroll
          for (int i = 0; i < N; ++i)
            if (x + i < xhi) dst[k][x + i] = tail[k][i];
        }
        /* x wrap of this row (xy / xz edges and corners of the destination): only the cells this tile owns */
        if (ext) dst[k][xeDst] = e[k];
      }
    }
  }
}

template <typename T> __device__ __noinline__ void wait_dirs(const PushParams<T> &Q, unsigned dirs, unsigned epoch) {
  if (threadIdx.x < kDirs && ((dirs >> threadIdx.x) & 1u)) {
    const unsigned *f = Q.myFlag[threadIdx.x];
    if (f) {
      const unsigned long long t0 = globaltimer_ns();
      while (int(ld_acquire_sys(f) - (epoch - 1)) < 0) {
        __nanosleep(64);
        if (globaltimer_ns() - t0 > Q.timeoutNs) {
          atomicCAS(Q.err, 0u, 200u + threadIdx.x);
          break;
        }
      }
    }
  }
}

/* thread 0 of a tile that stored into another GPU (after a CTA barrier) */
template <typename T> __device__ __noinline__ void signal_dirs(const PushParams<T> &Q, unsigned dirs, unsigned epoch) {
  __threadfence();
  unsigned ready = 0;
  for (int d = 0; d < kDirs; ++d)
    if ((dirs >> d) & 1u) {
      if (atomicAdd(Q.done + d, 1u) == Q.nDir[d] - 1) {
        Q.done[d] = 0;
        ready |= 1u << d;
      }
    }
  if (ready) {
    __threadfence_system();
    for (int d = 0; d < kDirs; ++d)
      if ((ready >> d) & 1u) st_flag_sys(Q.nbFlag[d], epoch);
  }
}

/* which push directions does a tile with rows [ylo, ylo+tr) and planes [z0, z1) own cells for? */
inline unsigned tile_dirs(int ylo, int tr, int z0, int z1, int sy, int sz, unsigned dirMask) {
  const bool ym = ylo < kR, yp = ylo + tr > sy - kR && ylo < sy, zm = z0 < kR, zp = z1 > sz - kR;
  unsigned f = 0;
  if (ym) f |= 1u << kYM;
  if (yp) f |= 1u << kYP;
  if (zm) f |= 1u << kZM;
  if (zp) f |= 1u << kZP;
  if (ym && zm) f |= 1u << kYMZM;
  if (ym && zp) f |= 1u << kYMZP;
  if (yp && zm) f |= 1u << kYPZM;
  if (yp && zp) f |= 1u << kYPZP;
  return f & dirMask;
}

struct ATile {
  int x0, y0;           // first cell / row of the tile (compute coordinates)
  int z0;               // first plane of the tile's z chunk
  unsigned short nz;    // planes of the chunk (chunks differ in depth: see astaroth_layers)
  unsigned char field, dirs, narrow, down; // down: march z downwards (neighbouring chunks alternate, see astaroth_tile)
  unsigned char pad[2];
};

template <typename T> struct TmaParams {
  T *dst[kMaxFields];      // compute cell (0,0,0) of next, per field
  const CUtensorMap *maps; // device array over the CURRENT curr buffers: [2 * field + narrow]
  const ATile *tiles;
  long long pitchX, pitchXY;
  int lox, loy, loz, hix, hiy, hiz; // cells to update, compute coordinates, half open
  int tileZ;
  int x0pad;               // index of compute x = 0 inside a padded row (TMA coordinates start at the row start)
  const PushParams<T> *push; // device pointer, nullptr: no pushes
  unsigned pushFlagMask;     // copy of push->flagMask (decided before anything is loaded)
  unsigned epoch;            // iteration being computed (1-based)
  int wrapX;               // > 0: the subdomain is its own +-x neighbour and is wrapX cells wide; the kernel then also
                           // writes its first / last three cells of every row into the opposite x halo of `next`
  T dt, eta;               // eta: weight of the cross terms (CROSS instantiations)
};

template <typename T, int LW, bool CROSS>
__device__ __forceinline__ void astaroth_tile(unsigned char *smem, const TmaParams<T> &P, const ATile &t) {
  typedef TmaGeom<T, LW> G;
  typedef typename Vec<T>::type V;
  constexpr int N = G::N;
  unsigned long long *full = reinterpret_cast<unsigned long long *>(smem + G::barOffset);
  unsigned long long *empty = full + kTStages;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, field = t.field;
  const int xl = lane & (LW - 1), sub = lane / LW;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTStages; ++s) {
      tma::mbar_init(&full[s], 1);
      tma::mbar_init(&empty[s], kTThreads / 32);
    }
    tma::mbar_fence_init();
  }
  const int z0 = t.z0, z1 = min(t.z0 + int(t.nz), P.hiz);
  /* tiles that exchange with another GPU first wait for that GPU's flags of the previous iteration (they are the only
     tiles that read halo cells it wrote); the TMA reads that follow are async-proxy accesses, ordered after the
     acquire by an explicit proxy fence of the issuing thread */
  const unsigned dirs = P.push ? unsigned(t.dirs) : 0u;
  if (dirs & P.pushFlagMask) wait_dirs<T>(*P.push, dirs & P.pushFlagMask, P.epoch);
  __syncthreads();
  if ((dirs & P.pushFlagMask) && threadIdx.x == 0) asm volatile("fence.proxy.async.global;" ::: "memory");

  const int nplanes = z1 - z0 + 2 * kR; // input planes z0-3 .. z1+2
  /* Neighbouring z chunks march in opposite directions, so the six planes both of them read are fetched by both at the
     same end of their march, at about the same time, and the second fetch hits L2 (marching everything upwards re-read
     them a whole tile duration apart: a third more DRAM reads than the minimum). The stencil is symmetric in z; only the
     z-coupled cross terms change sign. */
  const bool down = t.down != 0;
  const CUtensorMap *map = P.maps + 2 * field + (LW == 32 ? 0 : 1);
  const int cx = P.x0pad + t.x0 - kTMargin; // raw x of smem column 0
  const int cy = t.y0;                      // raw row of compute row y0 - 3 (raw = compute + 3)
  auto issue = [&](int j) {                 // input plane j = compute plane z0 - 3 + j = raw plane z0 + j
    const int sj = j % kTStages;
    if (j >= kTStages) tma::mbar_wait(&empty[sj], ((j / kTStages) - 1) & 1);
    tma::mbar_arrive_expect_tx(&full[sj], G::planeBytes);
    tma::load_3d(smem + sj * G::stageBytes, map, cx, cy, down ? z1 + 2 * kR - 1 - j : z0 + j, &full[sj]);
  };
  if (threadIdx.x == 0)
    for (int j = 0; j < kTStages - kTLag && j < nplanes; ++j) issue(j);

  const int r0 = 2 * (warp * G::kSub + sub);
  const int x0 = t.x0 + xl * N, y0 = t.y0 + r0;
  /* where the two rows of this lane go, decided once */
  T *dp[2];
  int mode[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int y = y0 + r;
    const bool in = y >= P.loy && y < P.hiy && x0 < P.hix && x0 + N > P.lox;
    mode[r] = !in ? 0 : ((x0 >= P.lox && x0 + N <= P.hix) ? 1 : 2);
    dp[r] = P.dst[field] + (long long)(down ? z1 - 1 : z0) * P.pitchXY + (long long)y * P.pitchX + x0;
  }
  /* Fused x faces (periodic wrap onto this subdomain): the lanes that own cells 0..2 or sx-3..sx-1 store their output
     vector a second time, sx cells further / earlier in the same row — that is where the opposite halo lives. The extra
     cells of the vector land in row padding. fp32 vectors are half a 32-byte sector: the other half (padding only) is
     written as well so that no sector is left partially written. */
  int qdelta = 0, qwide = 0;
  const bool qvec = P.wrapX % N == 0;
  if (P.wrapX > 0) {
    if (x0 < kR) {
      qdelta = P.wrapX;
      qwide = (N == 4 && x0 == 0 && qvec) ? N : 0;
    } else if (x0 + N > P.wrapX - kR && x0 < P.wrapX) {
      qdelta = -P.wrapX;
      qwide = (N == 4 && x0 + N == P.wrapX && qvec) ? -N : 0;
    }
  }
  const T c1 = T(1.5), c2 = T(-0.15), c3 = T(1.0 / 90.0), c0x3 = T(3) * T(-49.0 / 18.0);
  const T k1 = P.eta * CrossW<T>::m1(), k2 = P.eta * CrossW<T>::m2(), k3 = P.eta * CrossW<T>::m3();
  const T zs = down ? T(-1) : T(1), kz1 = zs * k1, kz2 = zs * k2, kz3 = zs * k3; // z-coupled cross terms follow the march
  const long long outStride = down ? -P.pitchXY : P.pitchXY;
  /* pending sums for the planes 1, 2, 3 behind the newest one, and the centres of those planes */
  T a1[2][N], a2[2][N], a3[2][N], h1[2][N], h2[2][N], h3[2][N];
  /* CROSS: sums collected AHEAD of the planes 1, 2, 3 in front of the newest one (the "below" halves of Dxz, Dyz, Dxyz) */
  T b1[2][N], b2[2][N], b3[2][N];
#pragma unroll
  for (int r = 0; r < 2; ++r)
#pragma unroll
    for (int i = 0; i < N; ++i) {
      a1[r][i] = a2[r][i] = a3[r][i] = h1[r][i] = h2[r][i] = h3[r][i] = T(0);
      b1[r][i] = b2[r][i] = b3[r][i] = T(0);
    }

  int s = 0;
  unsigned phase = 0;
  for (int j = 0; j < nplanes; ++j) {
    tma::mbar_wait(&full[s], phase);
    const T *pl = reinterpret_cast<const T *>(smem + s * G::stageBytes);
    /* smem row q holds compute row t.y0 + q - 3; smem column kTMargin + k holds compute x = t.x0 + k */
    const T *base = pl + (r0 + kR) * G::boxX + kTMargin + xl * N;
    T col[8][N]; // rows r0-3 .. r0+4 at my x: col[3], col[4] are my two centre rows
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const V v = *reinterpret_cast<const V *>(base + (q - 3) * G::boxX);
      memcpy(col[q], &v, sizeof(V));
    }
    const bool inPlane = j >= kR && j < nplanes - kR; // this plane is an output plane: it needs its in-plane sum
    T an[2][N];
    /* CROSS: e_k = [f(x+k) - f(x-k)] + [f(y+k) - f(y-k)] (+ the in-plane diagonal difference for k = 1) of THIS plane at
       my cells: what this plane contributes, with weight +-m_k, to the outputs k planes below / above it */
    T e1[2][N], e2[2][N], e3[2][N];
    if (CROSS || inPlane) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        /* x window of row r: cells x0 - kTMargin .. x0 + N + kTMargin - 1 */
        T win[N + 2 * kTMargin];
        const T *row = base + r * G::boxX;
#pragma unroll
        for (int v = 0; v < (N + 2 * kTMargin) / N; ++v) {
          if (v * N == kTMargin) {
#pragma unroll
            for (int i = 0; i < N; ++i) win[kTMargin + i] = col[3 + r][i];
          } else {
            const V tv = *reinterpret_cast<const V *>(row + v * N - kTMargin);
            memcpy(win + v * N, &tv, sizeof(V));
          }
        }
#pragma unroll
        for (int i = 0; i < N; ++i) {
          const int c = kTMargin + i;
          T lap = c0x3 * win[c];
          lap += c1 * (((win[c - 1] + win[c + 1]) + col[3 + r - 1][i]) + col[3 + r + 1][i]);
          lap += c2 * (((win[c - 2] + win[c + 2]) + col[3 + r - 2][i]) + col[3 + r + 2][i]);
          lap += c3 * (((win[c - 3] + win[c + 3]) + col[3 + r - 3][i]) + col[3 + r + 3][i]);
          /* contributions of the three planes below, whose centres are still in registers */
          lap += c1 * h1[r][i];
          lap += c2 * h2[r][i];
          lap += c3 * h3[r][i];
          an[r][i] = CROSS ? lap + b1[r][i] : lap;
          if (CROSS) {
            e1[r][i] = (win[c + 1] - win[c - 1]) + (col[3 + r + 1][i] - col[3 + r - 1][i]);
            e2[r][i] = (win[c + 2] - win[c - 2]) + (col[3 + r + 2][i] - col[3 + r - 2][i]);
            e3[r][i] = (win[c + 3] - win[c - 3]) + (col[3 + r + 3][i] - col[3 + r - 3][i]);
          }
        }
      }
      if (CROSS) {
        /* in-plane diagonals: row q of the bundle, k = |q - centre row| cells to either side, feeds Dxy of my two rows;
           the k = 1 difference is also the corner term's contribution of this plane and rides along in e1 */
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          T win[N + 2 * kTMargin];
          const T *row = base + (q - 3) * G::boxX;
#pragma unroll
          for (int v = 0; v < (N + 2 * kTMargin) / N; ++v) {
            if (v * N == kTMargin) {
#pragma unroll
              for (int i = 0; i < N; ++i) win[kTMargin + i] = col[q][i];
            } else {
              const V tv = *reinterpret_cast<const V *>(row + v * N - kTMargin);
              memcpy(win + v * N, &tv, sizeof(V));
            }
          }
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int dy = q - 3 - r;
            if (dy == 0 || dy < -kR || dy > kR) continue;
            const int k = dy < 0 ? -dy : dy;
            const T w = (k == 1 ? k1 : (k == 2 ? k2 : k3)) * T(dy < 0 ? -1 : 1);
#pragma unroll
            for (int i = 0; i < N; ++i) {
              const int c = kTMargin + i;
              const T d = win[c + k] - win[c - k];
              an[r][i] += w * d;
              if (k == 1) e1[r][i] += T(dy < 0 ? -1 : 1) * d;
            }
          }
        }
      }
    }
    if (!inPlane) {
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int i = 0; i < N; ++i) an[r][i] = T(0);
    }
    /* the slot can be refilled: everything needed from it is in registers */
    __syncwarp();
    if (lane == 0) tma::mbar_arrive(&empty[s]);
    if (threadIdx.x == 0 && j + kTStages - kTLag < nplanes) issue(j + kTStages - kTLag);
    if (++s == kTStages) {
      s = 0;
      phase ^= 1u;
    }
    /* this plane is "above" the three pending ones; the oldest of them is now complete */
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      T out[N];
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const T c = col[3 + r][i];
        if (CROSS) {
          out[i] = h3[r][i] + P.dt * ((a3[r][i] + c3 * c) + kz3 * e3[r][i]);
          a3[r][i] = (a2[r][i] + c2 * c) + kz2 * e2[r][i];
          a2[r][i] = (a1[r][i] + c1 * c) + kz1 * e1[r][i];
          b1[r][i] = b2[r][i] - kz1 * e1[r][i];
          b2[r][i] = b3[r][i] - kz2 * e2[r][i];
          b3[r][i] = -(kz3 * e3[r][i]);
        } else {
          out[i] = h3[r][i] + P.dt * (a3[r][i] + c3 * c);
          a3[r][i] = a2[r][i] + c2 * c;
          a2[r][i] = a1[r][i] + c1 * c;
        }
        a1[r][i] = an[r][i];
        h3[r][i] = h2[r][i];
        h2[r][i] = h1[r][i];
        h1[r][i] = c;
      }
      if (j >= 2 * kR) { // output plane z0 + j - 6 (marching down: z1 - 1 - (j - 6))
        if (mode[r] == 1) {
          V v;
          memcpy(&v, out, sizeof(V));
          *reinterpret_cast<V *>(dp[r]) = v;
          if (qdelta) {
            if (qvec) {
              *reinterpret_cast<V *>(dp[r] + qdelta) = v;
              if (qwide) *reinterpret_cast<V *>(dp[r] + qdelta + qwide) = v; // padding cells: any value will do
            } else { // row length not a multiple of the vector: element stores, face cells only
#pragma unroll
              for (int i = 0; i < N; ++i)
                if (x0 + i < kR || x0 + i >= P.wrapX - kR) dp[r][qdelta + i] = out[i];
            }
          }
        } else if (mode[r] == 2) {
#pragma unroll
          for (int i = 0; i < N; ++i)
            if (x0 + i >= P.lox && x0 + i < P.hix) {
              dp[r][i] = out[i];
              if (qdelta && (x0 + i < kR || x0 + i >= P.wrapX - kR)) dp[r][qdelta + i] = out[i];
            }
        }
        dp[r] += outStride;
      }
    }
  }
  if (dirs) {
    /* The plane loop above is untouched by the fusion. A boundary tile copies the rows it has just written (still in
       L2) into the neighbours' halos here, whole CTA, one warp per row. */
    __syncthreads();
    push_tile_dirs<T>(*P.push, field, dirs, t.x0, G::tileX, t.y0, G::tileY, z0, z1, P.dst[field], P.pitchX, P.pitchXY);
    if (dirs & P.pushFlagMask) {
      __syncthreads();
      if (threadIdx.x == 0) signal_dirs<T>(*P.push, dirs & P.pushFlagMask, P.epoch);
    }
  }
}

template <typename T, bool CROSS>
__global__ void __launch_bounds__(kTThreads, CROSS ? 1 : 2) astaroth_tma_kernel(const __grid_constant__ TmaParams<T> P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const ATile t = P.tiles[blockIdx.x];
  if (t.narrow)
    astaroth_tile<T, 8, CROSS>(smem, P, t);
  else
    astaroth_tile<T, 32, CROSS>(smem, P, t);
}

template <typename T>
__global__ void astaroth_init_kernel(T *logical0, long long pitchX, long long pitchXY, Dim3 raw, Dim3 haloOrigin, double period) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= raw.x || y >= raw.y) return;
  for (int z = blockIdx.z; z < raw.z; z += gridDim.z) {
    const bool inside = x >= kR && y >= kR && z >= kR && x < raw.x - kR && y < raw.y - kR && z < raw.z - kR;
    T v = T(-10);
    if (inside) {
      const double w = 2 * 3.14159 / period;
      v = T(sin(w * double(haloOrigin.x + x) + w * double(haloOrigin.y + y) + w * double(haloOrigin.z + z)));
    }
    logical0[(long long)z * pitchXY + (long long)y * pitchX + x] = v;
  }
}

} // namespace

struct AstarothSim::Impl {
  std::vector<RcStream> streams;
  std::vector<CUtensorMap *> maps[2]; // per subdomain: device array of two tensor maps (wide, narrow box) per field, per curr parity
  std::vector<Rect3> interior;
  std::vector<std::vector<Rect3>> exterior;
  /* tile tables of the TMA kernel, per subdomain: whole compute region (with push directions), interior, exterior slabs */
  struct Tiles {
    ATile *dev = nullptr;
    unsigned n = 0;
    bool hasNarrow = false; // decides how much shared memory a launch asks for
    int tileZ = 0;          // planes per z chunk of THIS table (balanced over the region's planes, see balanced_tile_z)
  };
  static Tiles upload(const std::vector<ATile> &t, int tileZ) {
    Tiles r;
    r.tileZ = tileZ;
    r.n = unsigned(t.size());
    for (const ATile &a : t) r.hasNarrow = r.hasNarrow || a.narrow;
    if (t.empty()) return r;
    CUDA_RUNTIME(cudaMalloc(&r.dev, t.size() * sizeof(ATile)));
    CUDA_RUNTIME(cudaMemcpy(r.dev, t.data(), t.size() * sizeof(ATile), cudaMemcpyHostToDevice));
    return r;
  }
  std::vector<Tiles> tilesAll, tilesInterior;
  std::vector<std::vector<Tiles>> tilesExterior;
  double exchangeSeconds = 0;
  /* free-running schedule: no host synchronisation inside the loop (exchange_async + events) */
  bool freeRun = false;
  /* the kernel writes the +-x halos of `next` itself (subdomains that wrap onto themselves in x) and exchange() leaves
     those faces out */
  bool fusedX = false;
  bool xHalosFromKernel = false; // the previous step's kernel filled the x halos of what is now `curr`
  /* y / z faces, yz edges and the x wrap of all of them pushed by the kernel too (with fusedX: no exchange() left in
     steady state, all 26 directions stay current) */
  bool fusedYZ = false;
  struct Push {
    void *dev[2] = {nullptr, nullptr}; // PushParams<T> per curr parity, in device memory
    unsigned dirMask = 0, flagMask = 0;
    unsigned nDir[kDirs] = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned *done = nullptr;
    std::vector<int> sameDevDeps; // local subdomains on the same GPU that push into me / I push into
    WaitSet finalWait;
  };
  std::vector<Push> push;
  unsigned iter = 0;       // iterations launched (the epoch published in the flags)
  unsigned *err = nullptr; // host-mapped error word for flag time-outs
  /* computeDone: end of each subdomain's kernel of the step being enqueued, computePrev: of the previous step (what
     same-GPU neighbours and the next exchange wait for); swapped every step so enqueueing threads never share one */
  std::vector<cudaEvent_t> computeDone, computePrev;
  std::vector<bool> computeRecorded; // computePrev[di] has been recorded
  detail::LaunchPool pool; // one enqueueing thread per local subdomain (1 process x N GPUs)
  bool launchThreads = true;
};

namespace {
const Dim3 kPushDir[kDirs] = {Dim3(0, -1, 0), Dim3(0, 1, 0), Dim3(0, 0, -1), Dim3(0, 0, 1), Dim3(0, -1, -1), Dim3(0, -1, 1), Dim3(0, 1, -1), Dim3(0, 1, 1)};

/* Tiles of the TMA kernel covering region [lo,hi) (compute coordinates) of a subdomain of size sz, for `fields` fields.
   vecN = cells per 16-byte vector (wide tile = 32 vecN cells x 16 rows, narrow = 8 vecN x 64). dirMask: push directions
   in use (0: no pushes); flagMask: those whose neighbour is another GPU; nDir[d] receives the number of CTAs that push
   to such a neighbour in direction d. Order: field-major, then (z, y, x), which keeps the tiles that share halo rows
   and planes resident together; tiles that push to another GPU come first (z directions, then the others) so that
   every direction is announced early in the kernel — see make_tiles() in src/jacobi.cu for the reasoning. */
/* Planes per z chunk for a region `planes` deep: as many chunks as the requested depth gives, all of (nearly) the same
   depth. 162 planes at depth 32 are 6 chunks of 27 instead of 5 x 32 + 2: the 2-plane chunk would load 8 planes for 2,
   and the three planes of the top face would straddle two chunks, which makes BOTH of them hand-shake with the
   neighbour GPU (half of all tiles instead of a third, each paying the push epilogue). */
int balanced_tile_z(int64_t planes, int tileZ) {
  if (planes <= 0) return tileZ;
  const int64_t chunks = (planes + tileZ - 1) / tileZ;
  return int(std::max<int64_t>(kR, (planes + chunks - 1) / chunks));
}

/* The z chunks of planes [lo, hi): balanced chunks of at most tileZ planes; with tailPlanes > 0 the top tailPlanes planes
   are cut into chunks of tailDepth instead — thin tiles that go to the end of the table and fill the kernel's ragged
   last wave in small pieces (same reasoning and measurements as z_layers() in src/jacobi.cu). */
struct ZChunks {
  std::vector<std::pair<int64_t, int64_t>> chunk;
  int depth = 0; // depth of the balanced chunks
};
ZChunks astaroth_layers(int64_t lo, int64_t hi, int tileZ, int tailPlanes = 0, int tailDepth = 8) {
  ZChunks r;
  const int64_t planes = hi - lo;
  r.depth = balanced_tile_z(planes, tileZ);
  if (planes <= 0) return r;
  tailDepth = std::max(kR, std::min(tailDepth, tileZ));
  int64_t tail = tailPlanes > 0 ? int64_t(tailPlanes) / tailDepth * tailDepth : 0;
  if (planes < tail + 2 * int64_t(tileZ)) tail = 0;
  const int64_t end = hi - tail;
  if (tail) r.depth = balanced_tile_z(end - lo, tileZ);
  for (int64_t z = lo; z < end; z += r.depth) r.chunk.emplace_back(z, std::min<int64_t>(z + r.depth, end));
  for (int64_t z = end; z < hi; z += tailDepth) r.chunk.emplace_back(z, std::min<int64_t>(z + tailDepth, hi));
  return r;
}

std::vector<ATile> make_atiles(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int fields, const ZChunks &zc, int vecN, unsigned dirMask, unsigned flagMask,
                               unsigned *nDir, bool narrowOk, bool flaggedFirst, bool zigzag) {
  std::vector<ATile> out;
  if (nDir)
    for (int d = 0; d < kDirs; ++d) nDir[d] = 0;
  if (hi.x <= lo.x || hi.y <= lo.y || hi.z <= lo.z) return out;
  const int64_t wideX = 32 * vecN, narrowX = 8 * vecN, narrowY = 4 * kTRows;
  std::vector<std::pair<int64_t, bool>> cols;
  for (int64_t x = lo.x / wideX * wideX; x < hi.x;) {
    const int64_t left = std::min<int64_t>(hi.x, sz.x) - x;
    if (narrowOk && left <= 3 * narrowX) {
      if (x + narrowX > lo.x) cols.emplace_back(x, true);
      x += narrowX;
    } else {
      cols.emplace_back(x, false);
      x += wideX;
    }
  }
  const int64_t nChunks = int64_t(zc.chunk.size());
  for (int f = 0; f < fields; ++f)
    for (int64_t c = 0; c < nChunks; ++c) {
      const int z0 = int(zc.chunk[size_t(c)].first), z1 = int(zc.chunk[size_t(c)].second);
      auto emit = [&](int64_t x0, int64_t y0, int64_t tr, bool nar) {
        ATile t;
        std::memset(&t, 0, sizeof(t));
        t.x0 = int(x0);
        t.y0 = int(y0);
        t.z0 = z0;
        t.nz = (unsigned short)(z1 - z0);
        t.field = (unsigned char)f;
        t.narrow = nar ? 1 : 0;
        t.down = (zigzag && (c & 1)) ? 1 : 0;
        t.dirs = (unsigned char)tile_dirs(int(y0), int(tr), z0, z1, int(sz.y), int(sz.z), dirMask);
        if (nDir)
          for (int d = 0; d < kDirs; ++d)
            if ((t.dirs & flagMask) >> d & 1u) ++nDir[d];
        out.push_back(t);
      };
      for (int64_t y0 = lo.y / kTRows * kTRows; y0 < hi.y; y0 += kTRows)
        for (const auto &col : cols)
          if (!col.second) emit(col.first, y0, kTRows, false);
      for (int64_t y0 = lo.y / narrowY * narrowY; y0 < hi.y; y0 += narrowY)
        for (const auto &col : cols)
          if (col.second) emit(col.first, y0, narrowY, true);
    }
  if (flaggedFirst) {
    /* tiles that push to another GPU first; tiles whose rows are mostly past the last row of the subdomain last (they
       take a fraction of a full tile and fill the kernel's last partial wave: see make_tiles() in src/jacobi.cu) */
    const unsigned zbits = (1u << kZM) | (1u << kZP);
    auto cheap = [&](const ATile &t) {
      const int64_t r = t.narrow ? narrowY : kTRows;
      return 2 * (std::min<int64_t>(t.y0 + r, hi.y) - std::max<int64_t>(t.y0, lo.y)) <= r || 2 * int64_t(t.nz) <= zc.depth;
    };
    auto prio = [&](const ATile &t) { return (t.dirs & flagMask & zbits) ? 0 : ((t.dirs & flagMask) ? 1 : (cheap(t) ? 3 : 2)); };
    std::stable_sort(out.begin(), out.end(), [&](const ATile &a, const ATile &b) { return prio(a) < prio(b); });
  }
  return out;
}

/* one PushParams<T> (device memory) for local subdomain di with `curr` = buffer `parity` */
template <typename T>
void *build_push(DistributedDomain &dd, size_t di, int parity, int fields, const NeighborInfo (*nbs)[kMaxFields], unsigned dirMask, unsigned flagMask,
                 const unsigned *nDir, unsigned *done, unsigned *err) {
  LocalDomain &ld = dd.domains()[di];
  PushParams<T> Q;
  std::memset(&Q, 0, sizeof(Q));
  const Dim3 sz = ld.size();
  for (int d = 0; d < kDirs; ++d) {
    if (!((dirMask >> d) & 1u)) continue;
    const Dim3 dir = kPushDir[d];
    for (int q = 0; q < fields; ++q) {
      const NeighborInfo &n = nbs[d][q];
      const long long npx = n.view.pitchY / (long long)sizeof(T), npxy = n.view.pitchZ / (long long)sizeof(T);
      T *interior = reinterpret_cast<T *>(n.view.base[parity ^ 1]) + (long long)kR * (npxy + npx + 1); // its NEXT buffer
      /* my first rows / planes land behind its last ones (+ halo), my last rows / planes before its first ones (- halo) */
      long long shift = 0;
      if (dir.y < 0) shift += n.size.y * npx;
      if (dir.y > 0) shift -= sz.y * npx;
      if (dir.z < 0) shift += n.size.z * npxy;
      if (dir.z > 0) shift -= sz.z * npxy;
      Q.nb[d][q] = interior + shift;
      Q.nbPitchX[d] = npx;
      Q.nbPitchXY[d] = npxy;
    }
    if ((flagMask >> d) & 1u) {
      /* I raise the slot "from direction -dir" in the neighbour; the neighbour raises slot "from dir" in me */
      Q.nbFlag[d] = nbs[d][0].control + kFusedFlagBase + DirectionMap<int>::slot(dir * -1);
      Q.myFlag[d] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(dir);
    }
    Q.nDir[d] = nDir[d];
  }
  Q.dirMask = dirMask;
  Q.flagMask = flagMask;
  Q.done = done;
  Q.err = err;
  Q.timeoutNs = flag_timeout_ns();
  Q.sx = int(sz.x);
  Q.sy = int(sz.y);
  Q.sz = int(sz.z);
  void *d = nullptr;
  CUDA_RUNTIME(cudaMalloc(&d, sizeof(Q)));
  CUDA_RUNTIME(cudaMemcpy(d, &Q, sizeof(Q), cudaMemcpyHostToDevice));
  return d;
}
} // namespace

static bool is_star(AstarothSim::Stencil st) { return st != AstarothSim::Stencil::Average7; }

namespace detail {
std::vector<TileBox> astaroth_tile_table(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int fields, int tileZ, int vecN, unsigned dirMask,
                                         unsigned flagMask, bool narrow, bool flaggedFirst, unsigned nDir[8], int *tileZUsed, int tailPlanes, int tailDepth) {
  const ZChunks zc = astaroth_layers(lo.z, hi.z, tileZ, tailPlanes, tailDepth);
  if (tileZUsed) *tileZUsed = zc.depth;
  std::vector<TileBox> out;
  for (const ATile &t : make_atiles(lo, hi, sz, fields, zc, vecN, dirMask, flagMask, nDir, narrow, flaggedFirst, true)) {
    const int64_t w = (t.narrow ? 8 : 32) * vecN, r = t.narrow ? 4 * kTRows : kTRows;
    TileBox b;
    b.lo = Dim3(std::max<int64_t>(t.x0, lo.x), std::max<int64_t>(t.y0, lo.y), t.z0);
    b.hi = Dim3(std::min<int64_t>(t.x0 + w, hi.x), std::min<int64_t>(t.y0 + r, hi.y), t.z0 + int64_t(t.nz));
    b.faces = t.dirs;
    b.shape = t.narrow;
    b.field = t.field;
    out.push_back(b);
  }
  return out;
}
} // namespace detail

AstarothSim::AstarothSim(size_t x, size_t y, size_t z, const Options &opt) : opt_(opt) {
  if (opt_.tileZ > 32768) opt_.tileZ = 32768; // ATile::nz is 16 bits wide
  if (opt_.fields < 1 || opt_.fields > kMaxFields) LOG_FATAL("AstarothSim supports 1.." << kMaxFields << " fields");
  dd_.reset(new DistributedDomain(x, y, z));
  if (!opt_.gpus.empty()) dd_->set_gpus(opt_.gpus);
  dd_->set_radius(kR);
  dd_->set_methods(opt_.methods);
  dd_->set_placement(opt_.placement);
  dd_->set_plan_file(opt_.planFile);
  for (int f = 0; f < opt_.fields; ++f) dd_->add_data(opt_.fp64 ? sizeof(double) : sizeof(float), "d" + std::to_string(f));
  dd_->realize();
  impl_.reset(new Impl);
  for (LocalDomain &ld : dd_->domains()) impl_->streams.emplace_back(ld.gpu());
  impl_->interior = dd_->get_interior();
  impl_->exterior = dd_->get_exterior();
  if (std::getenv("STENCIL_ASTAROTH_NO_TMA")) opt_.useTma = false;
  if (std::getenv("STENCIL_ASTAROTH_NO_NARROW")) opt_.narrowTiles = false;
  if (std::getenv("STENCIL_ASTAROTH_NO_ZIGZAG")) opt_.zigzag = false;
  if (std::getenv("STENCIL_ASTAROTH_NATURAL_ORDER")) opt_.flaggedFirst = false;
  if (const char *e = std::getenv("STENCIL_ASTAROTH_TAIL_PLANES")) opt_.tailPlanes = std::atoi(e);
  if (const char *e = std::getenv("STENCIL_ASTAROTH_TAIL_DEPTH")) opt_.tailDepth = std::atoi(e);
  if (const char *e = std::getenv("STENCIL_LAUNCH_THREADS")) impl_->launchThreads = std::atoi(e) != 0;
  for (LocalDomain &ld : dd_->domains()) {
    DevGuard g(ld.gpu());
    cudaEvent_t e;
    CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    impl_->computeDone.push_back(e);
    CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    impl_->computePrev.push_back(e);
    impl_->computeRecorded.push_back(false);
  }
  const bool star = is_star(opt_.stencil);
  if (opt_.autoSchedule && opt_.useTma && star) opt_.overlap = false;
  impl_->freeRun = !opt_.overlap && opt_.freeRunning && dd_->exchange_device_only() && !std::getenv("STENCIL_ASTAROTH_HOST_SYNC");
  const int vecN = opt_.fp64 ? Vec<double>::n : Vec<float>::n;
  {
    /* whole-region TMA launches only, every subdomain its own x neighbour, rows that keep the second store aligned */
    const int tileX = 32 * vecN;
    bool ok = !opt_.overlap && opt_.fusedXFaces && opt_.useTma && star && dd_->subdomain_grid().x == 1 && dd_->exchange_x_faces_all_local() &&
              !std::getenv("STENCIL_ASTAROTH_NO_FUSED_X");
    for (LocalDomain &ld : dd_->domains()) ok = ok && ld.size().x >= tileX / 2 && ld.size().x >= 2 * kR + vecN && ld.size().y >= kTRows / 2;
    impl_->fusedX = ok;
  }
  setup_fused_faces();
  if (opt_.useTma && star) {
    /* two tensor maps per field and buffer: the padded buffer as (row pitch, rows, planes) with the wide and the narrow box */
    std::vector<LocalDomain> &doms = dd_->domains();
    impl_->tilesAll.resize(doms.size());
    impl_->tilesInterior.resize(doms.size());
    impl_->tilesExterior.resize(doms.size());
    for (size_t di = 0; di < doms.size(); ++di) {
      LocalDomain &ld = doms[di];
      DevGuard g(ld.gpu());
      const Dim3 raw = ld.raw_size();
      for (int parity = 0; parity < 2; ++parity) {
        std::vector<CUtensorMap> host;
        for (int f = 0; f < opt_.fields; ++f) {
          char *buf = static_cast<char *>(ld.slab()) + ld.buffer_start_offset(size_t(f), parity);
          const Dim3 pitch = ld.pitch(size_t(f));
          if (opt_.fp64) {
            host.push_back(tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, pitch.x, raw.y, raw.z, TmaGeom<double, 32>::boxX, TmaGeom<double, 32>::boxY));
            host.push_back(tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, pitch.x, raw.y, raw.z, TmaGeom<double, 8>::boxX, TmaGeom<double, 8>::boxY));
          } else {
            host.push_back(tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, pitch.x, raw.y, raw.z, TmaGeom<float, 32>::boxX, TmaGeom<float, 32>::boxY));
            host.push_back(tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, pitch.x, raw.y, raw.z, TmaGeom<float, 8>::boxX, TmaGeom<float, 8>::boxY));
          }
        }
        CUtensorMap *d = nullptr;
        CUDA_RUNTIME(cudaMalloc(&d, host.size() * sizeof(CUtensorMap)));
        CUDA_RUNTIME(cudaMemcpy(d, host.data(), host.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
        impl_->maps[parity].push_back(d);
      }
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<double, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tma_smem_bytes<double>()));
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tma_smem_bytes<float>()));
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tma_smem_bytes<double>()));
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tma_smem_bytes<float>()));
      /* tile tables: the whole region carries the push directions (filled in by setup_fused_faces when fused) */
      const Dim3 sz = ld.size(), org = ld.origin();
      if (!impl_->tilesAll[di].dev)
      {
        const ZChunks zc = astaroth_layers(0, sz.z, opt_.tileZ, opt_.tailPlanes, opt_.tailDepth);
        impl_->tilesAll[di] = Impl::upload(make_atiles(Dim3(0, 0, 0), sz, sz, opt_.fields, zc, vecN, 0, 0, nullptr, opt_.narrowTiles, opt_.flaggedFirst, opt_.zigzag), zc.depth);
      }
      tma::keep_in_l2(impl_->streams[di], impl_->tilesAll[di].dev, size_t(impl_->tilesAll[di].n) * sizeof(ATile));
      {
        const ZChunks zc = astaroth_layers((impl_->interior[di].lo - org).z, (impl_->interior[di].hi - org).z, opt_.tileZ);
        impl_->tilesInterior[di] = Impl::upload(make_atiles(impl_->interior[di].lo - org, impl_->interior[di].hi - org, sz, opt_.fields, zc, vecN, 0, 0,
                                                            nullptr, opt_.narrowTiles, false, opt_.zigzag), zc.depth);
      }
      for (const Rect3 &r : impl_->exterior[di]) {
        const ZChunks zc = astaroth_layers((r.lo - org).z, (r.hi - org).z, opt_.tileZ);
        impl_->tilesExterior[di].push_back(
            Impl::upload(make_atiles(r.lo - org, r.hi - org, sz, opt_.fields, zc, vecN, 0, 0, nullptr, opt_.narrowTiles, false, opt_.zigzag), zc.depth));
      }
    }
  }
  init();
}

void AstarothSim::setup_fused_faces() {
  std::vector<LocalDomain> &doms = dd_->domains();
  impl_->push.resize(doms.size());
  impl_->tilesAll.resize(doms.size());
  bool ok = impl_->fusedX && impl_->freeRun && opt_.fusedYZFaces && !std::getenv("STENCIL_ASTAROTH_NO_FUSED_YZ");
  std::vector<std::vector<NeighborInfo>> all(doms.size()); // [di][d * kMaxFields + q]
  if (ok)
    for (size_t di = 0; di < doms.size() && ok; ++di) {
      all[di].resize(kDirs * kMaxFields);
      if (doms[di].size().y < 2 * kR || doms[di].size().z < 2 * kR || opt_.tileZ < kR) ok = false;
      for (int d = 0; d < kDirs && ok; ++d)
        for (int q = 0; q < opt_.fields; ++q) {
          const NeighborInfo n = dd_->neighbor(di, kPushDir[d], size_t(q));
          all[di][size_t(d * kMaxFields + q)] = n;
          /* every direction needs a neighbour we can store into; two processes on one GPU cannot hand-shake by spinning */
          if (!n.exists || !n.reachable || (n.sameDevice && !n.sameRank) || n.size.x != doms[di].size().x) ok = false;
        }
    }
  ok = dd_->comm().allreduce_sum(ok ? 0 : 1) == 0;
  impl_->fusedYZ = ok;
  if (!ok) return;
  CUDA_RUNTIME(cudaHostAlloc(&impl_->err, sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable));
  *impl_->err = 0;
  const int vecN = opt_.fp64 ? Vec<double>::n : Vec<float>::n;
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    DevGuard g(ld.gpu());
    Impl::Push &p = impl_->push[di];
    NeighborInfo nbs[kDirs][kMaxFields];
    std::set<int> deps;
    p.finalWait.n = 0;
    for (int d = 0; d < kDirs; ++d) {
      for (int q = 0; q < opt_.fields; ++q) nbs[d][q] = all[di][size_t(d * kMaxFields + q)];
      const NeighborInfo &nb = nbs[d][0];
      p.dirMask |= 1u << d;
      if (!(nb.sameRank && nb.sameDevice)) p.flagMask |= 1u << d;
      if (nb.sameRank && nb.sameDevice && nb.dom != int(di)) deps.insert(nb.dom);
    }
    p.sameDevDeps.assign(deps.begin(), deps.end());
    /* the tile table of the whole-region launch decides which CTAs take part in which hand-shake */
    const Dim3 sz = ld.size();
    const ZChunks zc = astaroth_layers(0, sz.z, opt_.tileZ, opt_.tailPlanes, opt_.tailDepth);
    impl_->tilesAll[di] = Impl::upload(make_atiles(Dim3(0, 0, 0), sz, sz, opt_.fields, zc, vecN, p.dirMask, p.flagMask, p.nDir, opt_.narrowTiles,
                                                   opt_.flaggedFirst, opt_.zigzag), zc.depth);
    CUDA_RUNTIME(cudaMalloc(&p.done, kDirs * sizeof(unsigned)));
    CUDA_RUNTIME(cudaMemset(p.done, 0, kDirs * sizeof(unsigned)));
    for (int parity = 0; parity < 2; ++parity)
      p.dev[parity] = opt_.fp64 ? build_push<double>(*dd_, di, parity, opt_.fields, nbs, p.dirMask, p.flagMask, p.nDir, p.done, impl_->err)
                                : build_push<float>(*dd_, di, parity, opt_.fields, nbs, p.dirMask, p.flagMask, p.nDir, p.done, impl_->err);
    for (int d = 0; d < kDirs; ++d)
      if ((p.flagMask >> d) & 1u) p.finalWait.flag[p.finalWait.n++] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(kPushDir[d]);
  }
}

AstarothSim::~AstarothSim() {
  if (dd_) sync();
  if (impl_ && dd_) {
    dd_->comm().barrier(); // nobody frees a slab that a neighbour's kernel may still be storing into
    for (size_t di = 0; di < impl_->push.size(); ++di) {
      DevGuard g(dd_->domains()[di].gpu());
      cudaFree(impl_->push[di].dev[0]);
      cudaFree(impl_->push[di].dev[1]);
      cudaFree(impl_->push[di].done);
    }
    for (size_t di = 0; di < dd_->domains().size(); ++di) {
      DevGuard g(dd_->domains()[di].gpu());
      if (di < impl_->tilesAll.size()) cudaFree(impl_->tilesAll[di].dev);
      if (di < impl_->tilesInterior.size()) cudaFree(impl_->tilesInterior[di].dev);
      if (di < impl_->tilesExterior.size())
        for (Impl::Tiles &t : impl_->tilesExterior[di]) cudaFree(t.dev);
    }
    if (impl_->err) cudaFreeHost(impl_->err);
  }
  if (impl_)
    for (cudaEvent_t e : impl_->computeDone) cudaEventDestroy(e);
  if (impl_)
    for (cudaEvent_t e : impl_->computePrev) cudaEventDestroy(e);
  if (impl_ && dd_)
    for (int parity = 0; parity < 2; ++parity)
      for (size_t di = 0; di < impl_->maps[parity].size(); ++di) {
        DevGuard g(dd_->domains()[di].gpu());
        cudaFree(impl_->maps[parity][di]);
      }
  impl_.reset();
  dd_.reset();
}

void AstarothSim::init() {
  if (impl_->iter > 0) sync(); // neighbours may still be storing the last step's faces into these buffers
  for (size_t di = 0; di < dd_->domains().size(); ++di) {
    LocalDomain &ld = dd_->domains()[di];
    DevGuard g(ld.gpu());
    const Dim3 raw = ld.raw_size();
    const dim3 block(64, 4, 1), grid(unsigned((raw.x + 63) / 64), unsigned((raw.y + 3) / 4), unsigned(std::min<int64_t>(raw.z, 64)));
    for (int f = 0; f < opt_.fields; ++f) {
      const Dim3 pitch = ld.pitch(size_t(f));
      if (opt_.fp64)
        astaroth_init_kernel<double><<<grid, block, 0, impl_->streams[di]>>>(static_cast<double *>(ld.curr_data(size_t(f))), pitch.x, pitch.x * pitch.y,
                                                                              raw, ld.halo_origin(), opt_.period);
      else
        astaroth_init_kernel<float><<<grid, block, 0, impl_->streams[di]>>>(static_cast<float *>(ld.curr_data(size_t(f))), pitch.x, pitch.x * pitch.y,
                                                                             raw, ld.halo_origin(), opt_.period);
      CUDA_RUNTIME(cudaGetLastError());
      ++launches_;
    }
  }
  sync();
  impl_->xHalosFromKernel = false; // curr was just rewritten: the next exchange carries the x faces again
}

template <typename T>
static void launch_region(AstarothSim::Stencil st, LocalDomain &ld, int fields, const Rect3 &reg, int tileZ, T dt, T eta, cudaStream_t stream) {
  if (reg.empty()) return;
  FieldParams<T> P;
  std::memset(&P, 0, sizeof(P));
  const Dim3 pitch = ld.pitch(0), org = ld.origin();
  const long long interiorOff = (long long)kR * (pitch.x * pitch.y + pitch.x + 1);
  for (int f = 0; f < fields; ++f) {
    P.src[f] = static_cast<const T *>(ld.curr_data(size_t(f))) + interiorOff;
    P.dst[f] = static_cast<T *>(ld.next_data(size_t(f))) + interiorOff;
  }
  P.pitchX = pitch.x;
  P.pitchXY = pitch.x * pitch.y;
  P.lox = int(reg.lo.x - org.x);
  P.loy = int(reg.lo.y - org.y);
  P.loz = int(reg.lo.z - org.z);
  P.hix = int(reg.hi.x - org.x);
  P.hiy = int(reg.hi.y - org.y);
  P.hiz = int(reg.hi.z - org.z);
  P.tileZ = tileZ;
  P.nChunks = (P.hiz - P.loz + tileZ - 1) / tileZ;
  P.dt = dt;
  constexpr int N = Vec<T>::n;
  const dim3 block(kLanes, kRows, 1);
  if (st == AstarothSim::Stencil::Star19Cross) {
    const dim3 grid(unsigned((P.hix - P.lox + kLanes - 1) / kLanes), unsigned((P.hiy - P.loy + kRows - 1) / kRows), unsigned(P.nChunks * fields));
    astaroth_cross_kernel<T><<<grid, block, 0, stream>>>(P, eta);
  } else {
    const int xSpan = P.hix - (P.lox / N) * N;
    const dim3 grid(unsigned((xSpan + kLanes * N - 1) / (kLanes * N)), unsigned((P.hiy - P.loy + kRows - 1) / kRows), unsigned(P.nChunks * fields));
    if (st == AstarothSim::Stencil::Star19)
      astaroth_kernel<T, true><<<grid, block, 0, stream>>>(P);
    else
      astaroth_kernel<T, false><<<grid, block, 0, stream>>>(P);
  }
  CUDA_RUNTIME(cudaGetLastError());
}

/* the TMA kernel on a region, through its tile table */
template <typename T>
static void launch_region_tma(bool cross, LocalDomain &ld, int fields, const CUtensorMap *maps, const ATile *tiles, unsigned nTiles, bool hasNarrow,
                              const Rect3 &reg,
                              int tileZ, T dt, T eta, cudaStream_t stream, bool wrapX = false, const void *push = nullptr, unsigned flagMask = 0,
                              unsigned epoch = 0) {
  if (!nTiles) return;
  TmaParams<T> P;
  std::memset(&P, 0, sizeof(P));
  const Dim3 pitch = ld.pitch(0), org = ld.origin();
  const long long interiorOff = (long long)kR * (pitch.x * pitch.y + pitch.x + 1);
  for (int f = 0; f < fields; ++f) P.dst[f] = static_cast<T *>(ld.next_data(size_t(f))) + interiorOff;
  P.maps = maps;
  P.tiles = tiles;
  P.pitchX = pitch.x;
  P.pitchXY = pitch.x * pitch.y;
  P.lox = int(reg.lo.x - org.x);
  P.loy = int(reg.lo.y - org.y);
  P.loz = int(reg.lo.z - org.z);
  P.hix = int(reg.hi.x - org.x);
  P.hiy = int(reg.hi.y - org.y);
  P.hiz = int(reg.hi.z - org.z);
  P.tileZ = tileZ;
  P.x0pad = int(ld.lead_elems(0)) + kR;
  P.dt = dt;
  P.eta = eta;
  P.wrapX = wrapX ? int(ld.size().x) : 0;
  P.push = static_cast<const PushParams<T> *>(push);
  P.pushFlagMask = flagMask;
  P.epoch = epoch;
  /* a table without narrow tiles needs only the wide tile's ring */
  const int smem = hasNarrow ? tma_smem_bytes<T>() : TmaGeom<T, 32>::smemBytes;
  if (cross)
    astaroth_tma_kernel<T, true><<<nTiles, kTThreads, smem, stream>>>(P);
  else
    astaroth_tma_kernel<T, false><<<nTiles, kTThreads, smem, stream>>>(P);
  CUDA_RUNTIME(cudaGetLastError());
}

void AstarothSim::step() {
  STENCIL_RANGE("AstarothSim::step");
  if (impl_->err && *impl_->err)
    LOG_FATAL("AstarothSim: timed out waiting for a neighbour GPU (code " << *impl_->err << ", iteration " << impl_->iter << ")");
  std::vector<LocalDomain> &doms = dd_->domains();
  const bool star = is_star(opt_.stencil), cross = opt_.stencil == Stencil::Star19Cross;
  /* tiles == nullptr: the region has no tile table (plain kernels only) */
  auto launch = [&](size_t di, const Rect3 &reg, const Impl::Tiles *tiles, bool whole) {
    if (reg.empty()) return;
    DevGuard g(doms[di].gpu());
    const Dim3 ext = reg.extent();
    const int tileX = 32 * (opt_.fp64 ? Vec<double>::n : Vec<float>::n);
    if (opt_.useTma && star && tiles && tiles->n && ext.x >= tileX / 2 && ext.y >= kTRows / 2) {
      const CUtensorMap *maps = impl_->maps[doms[di].parity()][di];
      const bool fused = whole && impl_->fusedYZ;
      const void *pushDev = fused ? impl_->push[di].dev[doms[di].parity()] : nullptr;
      const unsigned pushFlags = fused ? impl_->push[di].flagMask : 0u;
      if (opt_.fp64)
        launch_region_tma<double>(cross, doms[di], opt_.fields, maps, tiles->dev, tiles->n, tiles->hasNarrow, reg, tiles->tileZ, opt_.dt, opt_.crossWeight, impl_->streams[di],
                                  whole && impl_->fusedX, pushDev, pushFlags, impl_->iter);
      else
        launch_region_tma<float>(cross, doms[di], opt_.fields, maps, tiles->dev, tiles->n, tiles->hasNarrow, reg, tiles->tileZ, float(opt_.dt), float(opt_.crossWeight),
                                 impl_->streams[di], whole && impl_->fusedX, pushDev, pushFlags, impl_->iter);
      ++launches_;
      return;
    }
    if (opt_.fp64)
      launch_region<double>(opt_.stencil, doms[di], opt_.fields, reg, opt_.tileZ, opt_.dt, opt_.crossWeight, impl_->streams[di]);
    else
      launch_region<float>(opt_.stencil, doms[di], opt_.fields, reg, opt_.tileZ, float(opt_.dt), float(opt_.crossWeight), impl_->streams[di]);
    ++launches_;
  };
  auto tiles_all = [&](size_t di) { return di < impl_->tilesAll.size() ? &impl_->tilesAll[di] : nullptr; };
  /* x halos of `curr`: written by the previous step's kernel if there was one, else by this exchange */
  dd_->exchange_skip_local_x_faces(impl_->fusedX && impl_->xHalosFromKernel);
  const bool steady = impl_->fusedYZ && impl_->xHalosFromKernel; // every halo of `curr` was written by the previous kernels
  impl_->xHalosFromKernel = impl_->fusedX;
  ++impl_->iter;
  if (impl_->freeRun) {
    /* exchange_async() on the exchange streams, then one whole-region launch per subdomain ordered after it by an
       event; the next exchange is ordered after these launches the same way. The host only enqueues.
       With fused pushes there is no exchange at all after the first step: kernels of different GPUs hand-shake through
       flags, kernels of subdomains that share a GPU are ordered by events. */
    if (!steady) {
      for (size_t di = 0; di < doms.size(); ++di)
        if (impl_->computeRecorded[di]) dd_->exchange_after(di, impl_->computePrev[di]);
      dd_->exchange_async();
    }
    const std::function<void(size_t)> enqueue = [&](size_t di) {
      DevGuard g(doms[di].gpu());
      if (!steady) {
        CUDA_RUNTIME(cudaStreamWaitEvent(impl_->streams[di], dd_->exchange_done_event(di), 0));
      } else {
        for (int dep : impl_->push[di].sameDevDeps)
          if (impl_->computeRecorded[size_t(dep)]) CUDA_RUNTIME(cudaStreamWaitEvent(impl_->streams[di], impl_->computePrev[size_t(dep)], 0));
      }
      launch(di, doms[di].get_compute_region(), tiles_all(di), true);
      CUDA_RUNTIME(cudaEventRecord(impl_->computeDone[di], impl_->streams[di]));
    };
    if (doms.size() > 1 && impl_->launchThreads)
      impl_->pool.run(doms.size(), enqueue);
    else
      for (size_t di = 0; di < doms.size(); ++di) enqueue(di);
    impl_->computeDone.swap(impl_->computePrev);
    for (size_t di = 0; di < doms.size(); ++di) impl_->computeRecorded[di] = true;
    dd_->swap();
    return;
  }
  if (opt_.overlap)
    for (size_t di = 0; di < doms.size(); ++di) launch(di, impl_->interior[di], di < impl_->tilesInterior.size() ? &impl_->tilesInterior[di] : nullptr, false);
  const double t0 = Comm::wtime();
  dd_->exchange();
  impl_->exchangeSeconds += Comm::wtime() - t0;
  for (size_t di = 0; di < doms.size(); ++di) {
    if (opt_.overlap) {
      for (size_t si = 0; si < impl_->exterior[di].size(); ++si)
        launch(di, impl_->exterior[di][si], di < impl_->tilesExterior.size() && si < impl_->tilesExterior[di].size() ? &impl_->tilesExterior[di][si] : nullptr,
               false);
    } else {
      launch(di, doms[di].get_compute_region(), tiles_all(di), true);
    }
  }
  sync();
  dd_->swap();
}

void AstarothSim::sync() {
  const size_t nd = dd_->domains().size();
  const unsigned long long timeoutNs = flag_timeout_ns();
  const std::function<void(size_t)> drain = [&](size_t di) {
    DevGuard g(dd_->domains()[di].gpu());
    /* fused faces: my halos of the last iteration are complete only when every neighbour GPU raised its last flag */
    if (impl_->fusedYZ && impl_->iter > 0 && impl_->xHalosFromKernel)
      launch_wait_flags(impl_->push[di].finalWait, impl_->iter, impl_->err, timeoutNs, impl_->streams[di]);
    CUDA_RUNTIME(cudaStreamSynchronize(impl_->streams[di]));
  };
  if (nd > 1 && impl_->launchThreads)
    impl_->pool.run(nd, drain);
  else
    for (size_t di = 0; di < nd; ++di) drain(di);
  dd_->exchange_wait(); // no-op unless an asynchronous exchange is still in flight
  if (impl_->err && *impl_->err) LOG_FATAL("AstarothSim: timed out waiting for a neighbour GPU (code " << *impl_->err << ", iteration " << impl_->iter << ")");
}

double AstarothSim::run(int iters, double *exchangeSeconds, double *deviceSeconds) {
  sync();
  dd_->comm().barrier();
  impl_->exchangeSeconds = 0;
  const size_t nd = dd_->domains().size();
  std::vector<cudaEvent_t> e0(nd), e1(nd);
  for (size_t di = 0; di < nd; ++di) {
    DevGuard g(dd_->domains()[di].gpu());
    CUDA_RUNTIME(cudaEventCreate(&e0[di]));
    CUDA_RUNTIME(cudaEventCreate(&e1[di]));
    CUDA_RUNTIME(cudaEventRecord(e0[di], impl_->streams[di]));
  }
  const double t0 = Comm::wtime();
  for (int i = 0; i < iters; ++i) step();
  for (size_t di = 0; di < nd; ++di) {
    DevGuard g(dd_->domains()[di].gpu());
    CUDA_RUNTIME(cudaEventRecord(e1[di], impl_->streams[di]));
  }
  if (impl_->freeRun) sync();
  const double t = dd_->comm().allreduce_max(Comm::wtime() - t0);
  {
    float worst = 0;
    for (size_t di = 0; di < nd; ++di) {
      DevGuard g(dd_->domains()[di].gpu());
      CUDA_RUNTIME(cudaEventSynchronize(e1[di]));
      float ms = 0;
      CUDA_RUNTIME(cudaEventElapsedTime(&ms, e0[di], e1[di]));
      worst = std::max(worst, ms);
      cudaEventDestroy(e0[di]);
      cudaEventDestroy(e1[di]);
    }
    const double dev = dd_->comm().allreduce_max(double(worst) * 1e-3);
    if (deviceSeconds) *deviceSeconds = dev;
  }
  if (impl_->freeRun) {
    /* the loop never stops for the exchange, so its share is reported as the cost of the same exchange run alone */
    dd_->comm().barrier();
    const double e0 = Comm::wtime();
    dd_->exchange();
    impl_->exchangeSeconds = (Comm::wtime() - e0) * iters;
  }
  if (exchangeSeconds) *exchangeSeconds = dd_->comm().allreduce_max(impl_->exchangeSeconds);
  return t;
}

std::vector<double> AstarothSim::gather(int field) {
  sync();
  const Dim3 g = dd_->size();
  std::vector<double> out(size_t(g.flatten()), std::numeric_limits<double>::quiet_NaN());
  for (LocalDomain &ld : dd_->domains()) {
    const std::vector<unsigned char> b = ld.interior_to_host(size_t(field));
    const Dim3 sz = ld.size(), o = ld.origin();
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x, ++i) {
          const double v = opt_.fp64 ? reinterpret_cast<const double *>(b.data())[i] : double(reinterpret_cast<const float *>(b.data())[i]);
          out[size_t(((o.z + z) * g.y + o.y + y) * g.x + o.x + x)] = v;
        }
  }
  return out;
}

double astaroth_init_value(const Dim3 &p, double period, bool fp64) {
  const double w = 2 * 3.14159 / period;
  const double v = std::sin(w * double(p.x) + w * double(p.y) + w * double(p.z));
  return fp64 ? v : double(float(v));
}

template <typename T>
static void host_step(std::vector<double> &next, const std::vector<double> &curr, const Dim3 &size, AstarothSim::Stencil st, T dt, T eta) {
  const int64_t X = size.x, Y = size.y, Z = size.z;
  next.resize(curr.size());
  auto at = [&](int64_t x, int64_t y, int64_t z) { return T(curr[size_t((((z + 3 * Z) % Z) * Y + (y + 3 * Y) % Y) * X + (x + 3 * X) % X)]); };
  for (int64_t z = 0; z < Z; ++z)
    for (int64_t y = 0; y < Y; ++y)
      for (int64_t x = 0; x < X; ++x) {
        T v;
        if (st == AstarothSim::Stencil::Star19) {
          T xm[3], xp[3], ym[3], yp[3], zm[3], zp[3];
          for (int k = 0; k < 3; ++k) {
            xm[k] = at(x - k - 1, y, z);
            xp[k] = at(x + k + 1, y, z);
            ym[k] = at(x, y - k - 1, z);
            yp[k] = at(x, y + k + 1, z);
            zm[k] = at(x, y, z - k - 1);
            zp[k] = at(x, y, z + k + 1);
          }
          v = star19<T>(at(x, y, z), xm, xp, ym, yp, zm, zp, dt);
        } else if (st == AstarothSim::Stencil::Star19Cross) {
          T xm[3], xp[3], ym[3], yp[3], zm[3], zp[3];
          for (int k = 0; k < 3; ++k) {
            xm[k] = at(x - k - 1, y, z);
            xp[k] = at(x + k + 1, y, z);
            ym[k] = at(x, y - k - 1, z);
            yp[k] = at(x, y + k + 1, z);
            zm[k] = at(x, y, z - k - 1);
            zp[k] = at(x, y, z + k + 1);
          }
          auto rel = [&](int dx, int dy, int dz) { return at(x + dx, y + dy, z + dz); };
          v = star19<T>(at(x, y, z), xm, xp, ym, yp, zm, zp, dt) + dt * (eta * cross_terms<T>(rel));
        } else {
          v = average7<T>(at(x - 1, y, z), at(x, y - 1, z), at(x, y, z - 1), at(x + 1, y, z), at(x, y + 1, z), at(x, y, z + 1));
        }
        next[size_t((z * Y + y) * X + x)] = double(v);
      }
}

void astaroth_host_step(std::vector<double> &next, const std::vector<double> &curr, const Dim3 &size, AstarothSim::Stencil st, bool fp64, double dt,
                        double crossWeight) {
  if (fp64)
    host_step<double>(next, curr, size, st, dt, crossWeight);
  else
    host_step<float>(next, curr, size, st, float(dt), float(crossWeight));
}

} // namespace stencil
