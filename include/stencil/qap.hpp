#pragma once
/* Quadratic assignment: place n subdomains (weights w = bytes exchanged) on n GPUs (d = 1/bandwidth)
   minimising  cost(f) = sum_{a,b} w[a][b] * d[f[a]][f[b]]   with 0 * inf := 0.

   Parity: reference include/stencil/qap.hpp:15-172 —
     solve()        exact; lexicographic permutation scan from the identity keeping the FIRST strictly better one
     solve_catch()  steepest-descent pairwise-swap local search with O(n) incremental cost, strict improvement
   Tie-breaking (scan order + strict '<') matches so that placements — and the permutations pinned by
   test/test_cpu_qap.cpp:24-26,53-56,83-86 — are identical; the incremental update in solve_catch() applies its
   terms one by one in the same order (subtract old, swap, add new) so floating-point ties break the same way. */
#include <algorithm>
#include <cassert>
#include <cmath>
#include <numeric>
#include <vector>

#include "stencil/mat2d.hpp"

namespace qap {

namespace detail {

inline double cost_product(double we, double de) { return (we == 0 || de == 0) ? 0.0 : we * de; }

inline double cost(const Mat2D<double> &w, const Mat2D<double> &d, const std::vector<size_t> &f) {
  assert(w.shape().x == w.shape().y && d.shape() == w.shape() && w.shape().x == f.size());
  const size_t n = f.size();
  double c = 0;
  for (size_t a = 0; a < n; ++a)
    for (size_t b = 0; b < n; ++b) c += cost_product(w.at(a, b), d.at(f[a], f[b]));
  return c;
}

/* c += sign * (every term of cost(f) that involves position i or j), one term at a time */
inline void apply_touching(double &c, double sign, const Mat2D<double> &w, const Mat2D<double> &d,
                           const std::vector<size_t> &f, size_t i, size_t j) {
  const size_t n = f.size();
  for (size_t k = 0; k < n; ++k) {
    c += sign * cost_product(w.at(i, k), d.at(f[i], f[k]));
    c += sign * cost_product(w.at(j, k), d.at(f[j], f[k]));
    if (k != i && k != j) {
      c += sign * cost_product(w.at(k, i), d.at(f[k], f[i]));
      c += sign * cost_product(w.at(k, j), d.at(f[k], f[j]));
    }
  }
}

} // namespace detail

/* exact solver, O(n! * n^2) worst case; intended for n <= ~10 (GPUs of one node) */
inline std::vector<size_t> solve(const Mat2D<double> &w, const Mat2D<double> &d, double *costp = nullptr) {
  assert(w.shape() == d.shape() && w.shape().x == w.shape().y);
  const size_t n = w.shape().x;
  std::vector<size_t> f(n);
  std::iota(f.begin(), f.end(), size_t(0));
  std::vector<size_t> best = f;
  double bestCost = detail::cost(w, d, f);
  while (std::next_permutation(f.begin(), f.end())) {
    const double c = detail::cost(w, d, f);
    if (c < bestCost) {
      bestCost = c;
      best = f;
    }
  }
  if (costp) *costp = bestCost;
  return best;
}

/* 2-swap steepest descent ("CRAFT"-style); each sweep takes the best strictly-improving swap */
inline std::vector<size_t> solve_catch(const Mat2D<double> &w, const Mat2D<double> &d, double *costp = nullptr) {
  assert(w.shape() == d.shape() && w.shape().x == w.shape().y);
  const size_t n = w.shape().x;
  std::vector<size_t> best(n);
  std::iota(best.begin(), best.end(), size_t(0));
  double bestCost = detail::cost(w, d, best);
  for (;;) {
    bool improved = false;
    size_t bi = 0, bj = 0;
    double sweepCost = bestCost;
    std::vector<size_t> f = best;
    for (size_t i = 0; i < n; ++i) {
      for (size_t j = i + 1; j < n; ++j) {
        double c = bestCost;
        detail::apply_touching(c, -1.0, w, d, f, i, j);
        std::swap(f[i], f[j]);
        detail::apply_touching(c, +1.0, w, d, f, i, j);
        std::swap(f[i], f[j]);
        if (c < sweepCost) {
          sweepCost = c;
          bi = i;
          bj = j;
          improved = true;
        }
      }
    }
    if (!improved) break;
    /* The incremental estimate carries rounding error; with tied costs a swap and its inverse can each look like an
       improvement and the search would never end (the reference has this latent loop). Accept the swap only if the
       exactly recomputed cost really went down — the cost is then strictly decreasing, so the loop terminates. */
    std::swap(best[bi], best[bj]);
    const double exact = detail::cost(w, d, best);
    if (!(exact < bestCost)) {
      std::swap(best[bi], best[bj]);
      break;
    }
    bestCost = exact;
  }
  if (costp) *costp = bestCost;
  return best;
}

} // namespace qap
