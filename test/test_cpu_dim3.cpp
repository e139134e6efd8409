/* Dim3 / Rect3 / DirectionMap basics (the reference has no dedicated test; its typos in max/!=/[] motivated one). */
#include "harness.hpp"
#include <sstream>
#include "stencil/dim3.hpp"
#include "stencil/direction_map.hpp"
#include "stencil/rect3.hpp"

TEST_CASE("dim3 arithmetic and ordering", "[dim3]") {
  Dim3 a(1, 2, 3), b(4, 5, 6);
  REQUIRE(a + b == Dim3(5, 7, 9));
  REQUIRE(b - a == Dim3(3, 3, 3));
  REQUIRE(a * b == Dim3(4, 10, 18));
  REQUIRE(b / a == Dim3(4, 2, 2));
  REQUIRE(b % a == Dim3(0, 1, 0));
  REQUIRE(a * 2 == Dim3(2, 4, 6));
  REQUIRE(a * -1 == Dim3(-1, -2, -3));
  REQUIRE(a.flatten() == 6);
  REQUIRE(Dim3(3, 9, 4).max() == 9);
  REQUIRE(Dim3(3, 9, 4).min() == 3);
  REQUIRE(a != Dim3(1, 2, 4));
  REQUIRE(!(a != Dim3(1, 2, 3)));
  const Dim3 c(7, 8, 9);
  REQUIRE(c[0] == 7);
  REQUIRE(c[1] == 8);
  REQUIRE(c[2] == 9);
  /* x-major lexicographic */
  REQUIRE(Dim3(0, 9, 9) < Dim3(1, 0, 0));
  REQUIRE(Dim3(1, 0, 9) < Dim3(1, 1, 0));
  REQUIRE(Dim3(1, 1, 0) < Dim3(1, 1, 1));
  REQUIRE(!(Dim3(1, 1, 1) < Dim3(1, 1, 1)));
  REQUIRE(a.all_lt(4));
  REQUIRE(!a.all_lt(3));
  REQUIRE(a.any_gt(2));
  REQUIRE(a.all_ge(1));
}

TEST_CASE("dim3 wrap", "[dim3]") {
  REQUIRE(Dim3(-1, 0, 2).wrap(Dim3(2, 2, 2)) == Dim3(1, 0, 0));
  REQUIRE(Dim3(5, -3, 4).wrap(Dim3(3, 2, 4)) == Dim3(2, 1, 0));
  REQUIRE(Dim3(0, 0, 0).wrap(Dim3(1, 1, 1)) == Dim3(0, 0, 0));
  REQUIRE(Dim3(-1, 1, 0).wrapped(Dim3(1, 1, 1)) == Dim3(0, 0, 0));
}

TEST_CASE("dim3 make_block_dim", "[dim3]") {
  REQUIRE(Dim3::make_block_dim(Dim3(3, 512, 512), 512) == Dim3(4, 128, 1));
  REQUIRE(Dim3::make_block_dim(Dim3(512, 3, 512), 512) == Dim3(512, 1, 1));
  REQUIRE(Dim3::make_block_dim(Dim3(3, 3, 512), 512) == Dim3(4, 4, 32));
  REQUIRE(Dim3::make_block_dim(Dim3(3, 3, 3), 512) == Dim3(4, 4, 4));
  REQUIRE(Dim3::make_block_dim(Dim3(1, 512, 512), 512) == Dim3(1, 512, 1));
  REQUIRE(Dim3::make_block_dim(Dim3(1, 1, 4096), 4096) == Dim3(1, 1, 64));
  REQUIRE(nextPowerOfTwo(1) == 1);
  REQUIRE(nextPowerOfTwo(5) == 8);
  REQUIRE(nextPowerOfTwo(64) == 64);
}

TEST_CASE("dim3 and rect3 printing", "[dim3]") {
  std::ostringstream a, b;
  a << Dim3(1, -2, 3);
  REQUIRE(a.str() == "[1,-2,3]");
  b << Rect3(Dim3(0, 0, 0), Dim3(4, 5, 6));
  REQUIRE(b.str() == "[0,0,0]..<[4,5,6]");
  REQUIRE(Rect3(Dim3(1, 1, 1), Dim3(4, 5, 6)).extent() == Dim3(3, 4, 5));
}

TEST_CASE("direction map", "[dim3]") {
  DirectionMap<int> m;
  REQUIRE(m.at_dir(0, 0, 0) == 0);
  m.at_dir(-1, 0, 1) = 7;
  REQUIRE(m.at(0, 1, 2) == 7);
  REQUIRE(m.at_dir(Dim3(-1, 0, 1)) == 7);
  DirectionMap<int> n(3);
  REQUIRE(n.at_dir(1, 1, 1) == 3);
  REQUIRE(!(m == n));
  REQUIRE(DirectionMap<int>::slot(0, 0, 0) == 13);
  REQUIRE(DirectionMap<int>::slot(-1, -1, -1) == 0);
  REQUIRE(DirectionMap<int>::slot(1, 1, 1) == 26);
}
