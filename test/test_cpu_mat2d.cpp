/* Mat2D. Covers the behaviours exercised by the reference's test/test_cpu_mat2d.cpp (ctors, brace list,
   push_back, copy/move, row assignment) plus resize/reciprocal/permute. */
#include "harness.hpp"
#include <cmath>
#include "stencil/mat2d.hpp"

TEST_CASE("mat2d construction", "[mat2d]") {
  Mat2D<int> e;
  REQUIRE(e.shape() == Shape(0, 0));
  Mat2D<int> m(3, 2); // 3 columns, 2 rows
  REQUIRE(m.shape().x == 3);
  REQUIRE(m.shape().y == 2);
  Mat2D<int> f(3, 2, 7);
  REQUIRE(f.at(1, 2) == 7);
  Mat2D<int> b = {{1, 2, 3}, {4, 5, 6}};
  REQUIRE(b.shape() == Shape(3, 2));
  REQUIRE(b.at(0, 0) == 1);
  REQUIRE(b.at(1, 0) == 4);
  REQUIRE(b[1][2] == 6);
  REQUIRE(b[0][1] == 2);
}

TEST_CASE("mat2d push_back copy move rows", "[mat2d]") {
  Mat2D<double> m;
  m.push_back({1, 2});
  m.push_back({3, 4});
  REQUIRE(m.shape() == Shape(2, 2));
  REQUIRE(m.at(1, 1) == 4);
  Mat2D<double> c(m);
  REQUIRE(c == m);
  c.at(0, 0) = 9;
  REQUIRE(!(c == m));
  Mat2D<double> mv(std::move(c));
  REQUIRE(mv.at(0, 0) == 9);
  m[0] = m[1]; // row <- row
  REQUIRE(m.at(0, 0) == 3);
  REQUIRE(m.at(0, 1) == 4);
  m[1] = std::vector<double>{7, 8}; // row <- vector
  REQUIRE(m.at(1, 0) == 7);
  m /= 2.0;
  REQUIRE(m.at(1, 1) == 4);
}

TEST_CASE("mat2d resize reciprocal permute", "[mat2d]") {
  Mat2D<int> m = {{1, 2}, {3, 4}};
  m.resize(3, 3);
  REQUIRE(m.at(0, 0) == 1);
  REQUIRE(m.at(1, 1) == 4);
  REQUIRE(m.at(2, 2) == 0);
  m.resize(1, 1);
  REQUIRE(m.shape() == Shape(1, 1));
  REQUIRE(m.at(0, 0) == 1);
  Mat2D<double> bw = {{0, 2}, {4, 0}};
  Mat2D<double> d = make_reciprocal(bw);
  REQUIRE(std::isinf(d.at(0, 0)));
  REQUIRE(d.at(0, 1) == 0.5);
  REQUIRE(d.at(1, 0) == 0.25);
  Mat2D<int> p = {{0, 1, 2}, {10, 11, 12}, {20, 21, 22}};
  Mat2D<int> q = permute(p, {2, 0, 1});
  REQUIRE(q.at(0, 0) == 22);
  REQUIRE(q.at(0, 1) == 20);
  REQUIRE(q.at(1, 2) == 1);
}
