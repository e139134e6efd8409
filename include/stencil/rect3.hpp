#pragma once
/* Rect3: half-open axis-aligned box [lo, hi). Parity: reference include/stencil/rect3.hpp:13-27. */
#include <ostream>

#include "stencil/dim3.hpp"

class Rect3 {
public:
  Dim3 lo, hi;
  STENCIL_HD Rect3() {}
  STENCIL_HD Rect3(const Dim3 &_lo, const Dim3 &_hi) : lo(_lo), hi(_hi) {}
  STENCIL_HD Dim3 extent() const { return hi - lo; }
  STENCIL_HD bool empty() const { return hi.x <= lo.x || hi.y <= lo.y || hi.z <= lo.z; }
  STENCIL_HD bool contains(const Dim3 &p) const { return p.all_ge(lo) && p.all_lt(hi); }
  STENCIL_HD bool operator==(const Rect3 &r) const { return lo == r.lo && hi == r.hi; }
  STENCIL_HD bool operator!=(const Rect3 &r) const { return !(*this == r); }
};

inline std::ostream &operator<<(std::ostream &os, const Rect3 &r) { return os << r.lo << "..<" << r.hi; }
