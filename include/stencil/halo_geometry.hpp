#pragma once
/* Pure halo geometry shared by LocalDomain, the placement cost model and the exchange planner (no CUDA needed).
   Semantics follow the reference: include/stencil/local_domain.cuh:285-298 (halo_extent),
   src/local_domain.cu:56-95 (halo_pos). */
#include "stencil/dim3.hpp"
#include "stencil/radius.hpp"

namespace stencil {

/* extent of the halo region on side `dir` of a compute region of size sz: along an axis where dir is 0 the
   full size, else the radius on that side. dir = (0,0,0) gives sz. */
inline Dim3 halo_extent(const Dim3 &dir, const Dim3 &sz, const Radius &radius) {
  return Dim3(dir.x == 0 ? sz.x : int64_t(radius.x(int(dir.x))), dir.y == 0 ? sz.y : int64_t(radius.y(int(dir.y))),
              dir.z == 0 ? sz.z : int64_t(radius.z(int(dir.z))));
}

/* position (offset from the first allocated element, halo included) of the region on side `dir`:
   halo=true  -> the halo cells themselves (outside the compute region)
   halo=false -> the compute cells adjacent to that side (what a neighbour needs from us) */
inline Dim3 halo_pos(const Dim3 &dir, const Dim3 &sz, const Radius &radius, bool halo) {
  Dim3 p;
  const int64_t lo[3] = {int64_t(radius.x(-1)), int64_t(radius.y(-1)), int64_t(radius.z(-1))};
  for (int a = 0; a < 3; ++a) {
    const int64_t d = dir[a];
    if (d > 0)
      p[a] = sz[a] + (halo ? lo[a] : 0);
    else if (d < 0)
      p[a] = halo ? 0 : lo[a];
    else
      p[a] = lo[a];
  }
  return p;
}

} // namespace stencil
