/* bench-qap: time and solution cost of the two QAP solvers (CPU only).
   Output follows the reference's bin/bench_qap.cu:115-158: the reference's three tables (block-diagonal, random, matched) plus a permuted-match one with rows
     size CRAFT(s) cost exact(s) cost
   where the exact solver only runs for n < 9. Options (additions): --max-n (39), --reps (20). */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <limits>

#include "stencil/argparse.hpp"
#include "stencil/qap.hpp"

typedef std::chrono::steady_clock Clock;

static Mat2D<double> random_matrix(size_t n, int lo, int hi) {
  Mat2D<double> m(static_cast<int64_t>(n), static_cast<int64_t>(n));
  for (size_t i = 0; i < n; ++i)
    for (size_t j = 0; j < n; ++j) m.at(i, j) = (i == j) ? 0 : lo + std::rand() % (hi - lo + 1);
  return m;
}

/* fast links inside blocks of four, slow between blocks: the shape of a multi-socket NVLink machine */
static Mat2D<double> block_diagonal(size_t n) {
  Mat2D<double> m(static_cast<int64_t>(n), static_cast<int64_t>(n));
  for (size_t i = 0; i < n; ++i)
    for (size_t j = 0; j < n; ++j) m.at(i, j) = i == j ? 900 : (i / 4 == j / 4 ? 75 : 15);
  return m;
}

static void table(const char *title, size_t maxN, int reps, const std::function<void(size_t, Mat2D<double> &, Mat2D<double> &)> &make) {
  std::printf("%s\nsize CRAFT(s) cost exact(s) cost\n", title);
  for (size_t n = 2; n <= maxN; ++n) {
    Mat2D<double> w, bw;
    make(n, w, bw);
    Mat2D<double> d = make_reciprocal(bw);
    double cLocal = 0, cExact = std::numeric_limits<double>::quiet_NaN(), tLocal = 0, tExact = std::numeric_limits<double>::quiet_NaN();
    auto t0 = Clock::now();
    for (int r = 0; r < reps; ++r) qap::solve_catch(w, d, &cLocal);
    tLocal = std::chrono::duration<double>(Clock::now() - t0).count() / reps;
    if (n < 9) {
      t0 = Clock::now();
      const int er = n < 8 ? reps : 1;
      for (int r = 0; r < er; ++r) qap::solve(w, d, &cExact);
      tExact = std::chrono::duration<double>(Clock::now() - t0).count() / er;
    }
    std::printf("%zu %e %e %e %e\n", n, tLocal, cLocal, tExact, cExact);
  }
}

int main(int argc, char **argv) {
  size_t maxN = 39;
  int reps = 20;
  argparse::Parser p("QAP solver benchmark");
  p.add_option(maxN, "--max-n");
  p.add_option(reps, "--reps");
  if (!p.parse(argc, argv) || p.need_help()) {
    std::cout << p.help();
    return p.need_help() ? 0 : EXIT_FAILURE;
  }
  std::srand(1);
  table("block-diagonal bandwidth, random traffic", maxN, reps,
        [](size_t n, Mat2D<double> &w, Mat2D<double> &bw) {
          w = random_matrix(n, 0, 100);
          bw = block_diagonal(n);
        });
  table("random bandwidth, random traffic", maxN, reps, [](size_t n, Mat2D<double> &w, Mat2D<double> &bw) {
    w = random_matrix(n, 0, 100);
    bw = random_matrix(n, 1, 100);
    for (size_t i = 0; i < n; ++i) bw.at(i, i) = 900;
  });
  /* the reference's third family (bin/bench_qap.cu:32-48 there): the distance matrix is the elementwise reciprocal of the
     traffic, so the identity is already the perfect match and the local search stops after one sweep */
  table("matched: bandwidth equal to traffic (identity is optimal)", maxN, reps, [](size_t n, Mat2D<double> &w, Mat2D<double> &bw) {
    w = random_matrix(n, 1, 100);
    bw = w;
  });
  /* a harder fourth family: the match exists but is a permutation away, one swap per sweep */
  table("traffic matched to a permuted block-diagonal bandwidth", maxN, reps, [](size_t n, Mat2D<double> &w, Mat2D<double> &bw) {
    bw = block_diagonal(n);
    std::vector<size_t> perm(n);
    for (size_t i = 0; i < n; ++i) perm[i] = (i * 7 + 3) % n;
    bool ok = true; // fall back to a rotation when 7 is not coprime with n
    std::vector<char> seen(n, 0);
    for (size_t v : perm) {
      if (seen[v]) ok = false;
      seen[v] = 1;
    }
    if (!ok)
      for (size_t i = 0; i < n; ++i) perm[i] = (i + 1) % n;
    w = permute(bw, perm);
    for (size_t i = 0; i < n; ++i) w.at(i, i) = 0;
  });
  return 0;
}
