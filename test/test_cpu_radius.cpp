/* Radius. Mirrors the assertions of the reference's test/test_cpu_radius.cpp and adds face/edge/corner coverage. */
#include "harness.hpp"
#include "stencil/radius.hpp"

TEST_CASE("radius constant", "[radius]") {
  Radius r = Radius::constant(2);
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; ++x) REQUIRE(r.dir(x, y, z) == 2);
  REQUIRE(r.x(1) == 2);
  REQUIRE(r.y(-1) == 2);
  REQUIRE(r.z(0) == 2);
}

TEST_CASE("radius face edge corner", "[radius]") {
  Radius r = Radius::face_edge_corner(3, 2, 1);
  REQUIRE(r.dir(0, 0, 0) == 0);
  REQUIRE(r.dir(1, 0, 0) == 3);
  REQUIRE(r.dir(0, -1, 0) == 3);
  REQUIRE(r.dir(1, 1, 0) == 2);
  REQUIRE(r.dir(0, -1, 1) == 2);
  REQUIRE(r.dir(1, 1, 1) == 1);
  REQUIRE(r.dir(-1, 1, -1) == 1);
  REQUIRE(r.max() == 3);
  Radius s;
  REQUIRE(s.dir(1, 0, 0) == 0); // default-constructed radius is all zero
  s.set_face(3);
  s.set_edge(2);
  s.set_corner(1);
  REQUIRE(s == r);
  s.dir(Dim3(1, 0, 0)) = 5;
  REQUIRE(s != r);
  REQUIRE(s.x(1) == 5);
  REQUIRE(s.x(-1) == 3);
}
