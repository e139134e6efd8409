#pragma once
/* Statistics: running sample set with min/max/avg/stddev/median/trimean.
   Parity: reference bin/statistics.hpp:6-19, bin/statistics.cpp:7-55. trimean keeps the reference definition
   (x[n/4] + 2 x[2(n/4)] + x[3(n/4)]) / 4 on the sorted samples because benchmark CSV columns depend on it;
   med() of an even count is the mean of the two middle samples (the reference adds them and reads one past). */
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <numeric>
#include <vector>

class Statistics {
  std::vector<double> x_;

  std::vector<double> sorted() const {
    std::vector<double> s(x_);
    std::sort(s.begin(), s.end());
    return s;
  }

public:
  void clear() { x_.clear(); }
  void insert(double d) { x_.push_back(d); }
  size_t count() const noexcept { return x_.size(); }
  const std::vector<double> &samples() const noexcept { return x_; }

  double min() const { return x_.empty() ? std::nan("") : *std::min_element(x_.begin(), x_.end()); }
  double max() const { return x_.empty() ? std::nan("") : *std::max_element(x_.begin(), x_.end()); }
  double avg() const { return x_.empty() ? std::nan("") : std::accumulate(x_.begin(), x_.end(), 0.0) / double(x_.size()); }
  /* sample standard deviation (n-1) */
  double stddev() const {
    if (x_.size() < 2) return std::nan("");
    const double m = avg();
    double acc = 0;
    for (double v : x_) acc += (v - m) * (v - m);
    return std::sqrt(acc / double(x_.size() - 1));
  }
  double med() const {
    if (x_.empty()) return std::nan("");
    const std::vector<double> s = sorted();
    const size_t n = s.size();
    return (n % 2) ? s[n / 2] : 0.5 * (s[n / 2 - 1] + s[n / 2]);
  }
  double trimean() const {
    if (x_.empty()) return std::nan("");
    const std::vector<double> s = sorted();
    const size_t q = s.size() / 4;
    return (s[q] + 2 * s[2 * q] + s[3 * q]) / 4;
  }
};
