#pragma once
/* Boundary: per-direction "does the domain wrap?" flags (default: periodic everywhere).
   Parity: reference include/stencil/boundary.hpp:7-26 (declared there, used nowhere). Here
   DistributedDomain::set_boundary() consumes it: a non-wrapping direction simply plans no message across the
   global domain edge. */
#include "stencil/direction_map.hpp"

class Boundary {
  DirectionMap<bool> wraps_;

public:
  Boundary() : wraps_(true) {}
  bool &dir(int x, int y, int z) { return wraps_.at_dir(x, y, z); }
  const bool &dir(int x, int y, int z) const { return wraps_.at_dir(x, y, z); }
  void set_dir(int x, int y, int z, bool wraps) noexcept { wraps_.at_dir(x, y, z) = wraps; } // reference spelling (:18)
  static Boundary periodic() { return Boundary(); }
  static Boundary none() {
    Boundary b;
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x) b.dir(x, y, z) = false;
    return b;
  }
  bool operator==(const Boundary &o) const { return wraps_ == o.wraps_; }
};
