/* QAP solvers. Expected permutations are the ones the reference pins (test/test_cpu_qap.cpp:24-26,53-56,83-86):
   they fix the scan order and the strict-improvement tie-breaking of both solvers. */
#include "harness.hpp"
#include <cstdlib>
#include <limits>
#include "stencil/qap.hpp"

static const double inf = std::numeric_limits<double>::infinity();

TEST_CASE("qap unbalanced triangle", "[qap]") {
  Mat2D<double> bw = {{inf, 1, 10}, {1, inf, 1}, {10, 1, inf}}; // fast link 0-2
  Mat2D<double> comm = {{0, 10, 1}, {10, 0, 1}, {1, 1, 0}};     // heavy traffic 0-1
  Mat2D<double> dist = make_reciprocal(bw);
  double cost = -1;
  const std::vector<size_t> f = qap::solve(comm, dist, &cost);
  REQUIRE((f == std::vector<size_t>{0, 2, 1}));
  REQUIRE(cost == qap::detail::cost(comm, dist, f));
}

static Mat2D<double> p9_bw() { return {{900, 75, 64, 64}, {75, 900, 64, 64}, {64, 64, 900, 75}, {64, 64, 75, 900}}; }
static Mat2D<double> p9_comm() { return {{7, 5, 10, 1}, {5, 7, 1, 10}, {10, 1, 7, 5}, {1, 10, 5, 7}}; }

TEST_CASE("qap p9 exact", "[qap]") {
  Mat2D<double> dist = make_reciprocal(p9_bw());
  REQUIRE((qap::solve(p9_comm(), dist) == std::vector<size_t>{0, 2, 1, 3}));
}

TEST_CASE("qap p9 local search", "[qap]") {
  Mat2D<double> dist = make_reciprocal(p9_bw());
  double cExact = 0, cLocal = 0;
  qap::solve(p9_comm(), dist, &cExact);
  const std::vector<size_t> f = qap::solve_catch(p9_comm(), dist, &cLocal);
  REQUIRE((f == std::vector<size_t>{3, 1, 2, 0}));
  REQUIRE(std::abs(cLocal - cExact) < 1e-12); // same optimum, different representative
}

TEST_CASE("qap zero times infinity is zero", "[qap]") {
  Mat2D<double> w = {{0, 1}, {1, 0}};
  Mat2D<double> d = {{inf, 2}, {2, inf}};
  REQUIRE(qap::detail::cost(w, d, {0, 1}) == 4);
}

TEST_CASE("qap local search never worse than identity, random 64", "[qap]") {
  std::srand(7);
  Mat2D<double> bw(64, 64), comm(64, 64);
  for (size_t i = 0; i < 64; ++i)
    for (size_t j = 0; j < 64; ++j) {
      bw.at(i, j) = 1 + std::rand() % 1000;
      comm.at(i, j) = std::rand() % 1000;
    }
  Mat2D<double> dist = make_reciprocal(bw);
  std::vector<size_t> id(64);
  for (size_t i = 0; i < 64; ++i) id[i] = i;
  double c = 0;
  const std::vector<size_t> f = qap::solve_catch(comm, dist, &c);
  REQUIRE(c <= qap::detail::cost(comm, dist, id));
  REQUIRE(std::abs(c - qap::detail::cost(comm, dist, f)) < 1e-6 * c); // incremental cost stays consistent
  std::vector<size_t> sorted = f;
  std::sort(sorted.begin(), sorted.end());
  REQUIRE(sorted == id); // still a permutation
}

TEST_CASE("qap local search matches exact on small random instances", "[qap]") {
  std::srand(11);
  int matches = 0;
  for (int trial = 0; trial < 20; ++trial) {
    Mat2D<double> bw(6, 6), comm(6, 6);
    for (size_t i = 0; i < 6; ++i)
      for (size_t j = 0; j < 6; ++j) {
        bw.at(i, j) = 1 + std::rand() % 50;
        comm.at(i, j) = (i == j) ? 0 : std::rand() % 50;
      }
    Mat2D<double> dist = make_reciprocal(bw);
    double ce = 0, cl = 0;
    qap::solve(comm, dist, &ce);
    qap::solve_catch(comm, dist, &cl);
    REQUIRE(ce <= cl + 1e-9);
    if (cl <= ce * 1.05) ++matches;
  }
  REQUIRE(matches >= 10); // a local optimum within 5 % most of the time
}
