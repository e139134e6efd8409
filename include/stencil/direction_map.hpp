#pragma once
/* DirectionMap<T>: one T per direction vector in {-1,0,1}^3.
   Parity: reference include/stencil/direction_map.hpp:11-58. Stored flat (27 entries, x fastest) so it is a
   trivially-copyable POD usable in kernel parameters; default construction value-initialises. */
#include <cassert>

#include "stencil/dim3.hpp"

template <typename T> class DirectionMap {
  T data_[27];

public:
  typedef int index_type;
  STENCIL_HD DirectionMap() {
    for (int i = 0; i < 27; ++i) data_[i] = T();
  }
  STENCIL_HD explicit DirectionMap(const T &v) {
    for (int i = 0; i < 27; ++i) data_[i] = v;
  }
  /* index with each component in {0,1,2} */
  STENCIL_HD T &at(int x, int y, int z) { return data_[(z * 3 + y) * 3 + x]; }
  STENCIL_HD const T &at(int x, int y, int z) const { return data_[(z * 3 + y) * 3 + x]; }
  /* index with a direction vector, each component in {-1,0,1} */
  STENCIL_HD T &at_dir(int x, int y, int z) { return at(x + 1, y + 1, z + 1); }
  STENCIL_HD const T &at_dir(int x, int y, int z) const { return at(x + 1, y + 1, z + 1); }
  STENCIL_HD T &at_dir(const Dim3 &d) { return at_dir(int(d.x), int(d.y), int(d.z)); }
  STENCIL_HD const T &at_dir(const Dim3 &d) const { return at_dir(int(d.x), int(d.y), int(d.z)); }
  STENCIL_HD bool operator==(const DirectionMap &r) const {
    for (int i = 0; i < 27; ++i)
      if (!(data_[i] == r.data_[i])) return false;
    return true;
  }
  /* flat slot of a direction; slot 13 is (0,0,0) */
  STENCIL_HD static int slot(int x, int y, int z) { return ((z + 1) * 3 + (y + 1)) * 3 + (x + 1); }
  STENCIL_HD static int slot(const Dim3 &d) { return slot(int(d.x), int(d.y), int(d.z)); }
};
