#pragma once
/* Accessor<T>: (raw pointer, origin, pitch) view so kernels index a subdomain with GLOBAL coordinates.
   acc[p] = raw[(p - origin) linearised with pitch.x, pitch.y].
   Parity: reference include/stencil/accessor.hpp:13-44. Here pitch.x is the PADDED row pitch of LocalDomain
   (rows are 128 B aligned for TMA / 16 B vector access) — callers that only go through operator[] do not notice. */
#include <cassert>

#include "stencil/dim3.hpp"

template <typename T> class Accessor {
  T *raw_;
  Dim3 origin_; // coordinate of raw_[0] (includes the negative-side halo)
  Dim3 pitch_;  // elements; pitch_.x = row pitch, pitch_.y = rows per plane

public:
  STENCIL_HD Accessor() : raw_(nullptr) {}
  STENCIL_HD Accessor(T *raw, const Dim3 &origin, const Dim3 &pitch) : raw_(raw), origin_(origin), pitch_(pitch) {}

  STENCIL_HD STENCIL_INLINE T &operator[](const Dim3 &p) noexcept { return raw_[offset(p)]; }
  STENCIL_HD STENCIL_INLINE const T &operator[](const Dim3 &p) const noexcept { return raw_[offset(p)]; }
  STENCIL_HD STENCIL_INLINE T &at(int64_t x, int64_t y, int64_t z) noexcept {
    return raw_[((z - origin_.z) * pitch_.y + (y - origin_.y)) * pitch_.x + (x - origin_.x)];
  }
  STENCIL_HD STENCIL_INLINE const T &at(int64_t x, int64_t y, int64_t z) const noexcept {
    return raw_[((z - origin_.z) * pitch_.y + (y - origin_.y)) * pitch_.x + (x - origin_.x)];
  }

  STENCIL_HD const Dim3 &origin() const noexcept { return origin_; }
  STENCIL_HD const Dim3 &pitch() const noexcept { return pitch_; }
  STENCIL_HD T *raw() const noexcept { return raw_; }

private:
  STENCIL_HD STENCIL_INLINE int64_t offset(const Dim3 &p) const noexcept {
    const Dim3 o = p - origin_;
#ifndef NDEBUG
    assert(o.x >= 0 && o.y >= 0 && o.z >= 0);
#endif
    return (o.z * pitch_.y + o.y) * pitch_.x + o.x;
  }
};
