#pragma once
/* Radius: stencil radius per direction (27 entries). A halo for direction d exists on the receiver iff
   dir(d) > 0; allocation extents only use the six axis entries.
   Parity: reference include/stencil/radius.hpp:14-104. */
#include <cstddef>

#include "stencil/direction_map.hpp"

class Radius {
  DirectionMap<size_t> r_;

public:
  STENCIL_HD Radius() {}

  STENCIL_HD size_t &dir(int x, int y, int z) { return r_.at_dir(x, y, z); }
  STENCIL_HD const size_t &dir(int x, int y, int z) const { return r_.at_dir(x, y, z); }
  STENCIL_HD size_t &dir(const Dim3 &d) { return r_.at_dir(d); }
  STENCIL_HD const size_t &dir(const Dim3 &d) const { return r_.at_dir(d); }

  /* axis shortcuts: d in {-1,0,1} */
  STENCIL_HD const size_t &x(int d) const { return r_.at_dir(d, 0, 0); }
  STENCIL_HD const size_t &y(int d) const { return r_.at_dir(0, d, 0); }
  STENCIL_HD const size_t &z(int d) const { return r_.at_dir(0, 0, d); }

  STENCIL_HD bool operator==(const Radius &o) const { return r_ == o.r_; }
  STENCIL_HD bool operator!=(const Radius &o) const { return !(r_ == o.r_); }

  void set_face(size_t r) { set_by_order(1, r); }
  void set_edge(size_t r) { set_by_order(2, r); }
  void set_corner(size_t r) { set_by_order(3, r); }

  /* largest entry over all 26 directions */
  size_t max() const {
    size_t m = 0;
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x)
          if (x || y || z) m = dir(x, y, z) > m ? dir(x, y, z) : m;
    return m;
  }

  static Radius constant(size_t r) {
    Radius out;
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x) out.dir(x, y, z) = r;
    return out;
  }

  static Radius face_edge_corner(size_t face, size_t edge, size_t corner) {
    Radius out;
    out.set_face(face);
    out.set_edge(edge);
    out.set_corner(corner);
    out.dir(0, 0, 0) = 0;
    return out;
  }

private:
  /* order = number of non-zero components: 1 face, 2 edge, 3 corner */
  void set_by_order(int order, size_t r) {
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x)
          if ((x != 0) + (y != 0) + (z != 0) == order) dir(x, y, z) = r;
  }
};
