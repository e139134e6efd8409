#pragma once
/* Mat2D<T>: dense row-major matrix used for bandwidth / communication / distance matrices.
   Parity: reference include/stencil/mat2d.hpp:9-220 — ctor is (cols, rows[, value]); at(row, col);
   operator[] yields a row view; resize keeps the top-left block; push_back appends a row;
   make_reciprocal maps 0 -> +inf. permute() works here (the reference's version never compiled). */
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <initializer_list>
#include <limits>
#include <ostream>
#include <vector>

struct Shape {
  uint64_t x; // columns
  uint64_t y; // rows
  Shape(uint64_t _x, uint64_t _y) : x(_x), y(_y) {}
  uint64_t flatten() const noexcept { return x * y; }
  bool operator==(const Shape &r) const noexcept { return x == r.x && y == r.y; }
  bool operator!=(const Shape &r) const noexcept { return !(*this == r); }
};

template <typename T> class Mat2D {
  std::vector<T> data_;
  Shape shape_;

  template <typename P> class RowView {
    friend class Mat2D;
    P *p_;
    int64_t n_;
    RowView(P *p, int64_t n) : p_(p), n_(n) {}

  public:
    P &operator[](int64_t i) const noexcept { return p_[i]; }
    P *begin() const { return p_; }
    P *end() const { return p_ + n_; }
    int64_t size() const { return n_; }
    /* element-wise overwrite from another row or a vector (sizes must agree) */
    template <typename Q> const RowView &operator=(const RowView<Q> &r) const {
      assert(r.n_ == n_);
      std::copy(r.begin(), r.end(), p_);
      return *this;
    }
    const RowView &operator=(const RowView &r) const {
      assert(r.n_ == n_);
      std::copy(r.begin(), r.end(), p_);
      return *this;
    }
    const RowView &operator=(const std::vector<typename std::remove_const<P>::type> &v) const {
      assert(int64_t(v.size()) == n_);
      std::copy(v.begin(), v.end(), p_);
      return *this;
    }
  };

public:
  typedef RowView<T> Row;
  typedef RowView<const T> ConstRow;

  Mat2D() : shape_(0, 0) {}
  Mat2D(int64_t cols, int64_t rows) : data_(size_t(cols * rows)), shape_(cols, rows) {}
  Mat2D(int64_t cols, int64_t rows, const T &v) : data_(size_t(cols * rows), v), shape_(cols, rows) {}
  explicit Mat2D(Shape s) : Mat2D(int64_t(s.x), int64_t(s.y)) {}
  Mat2D(Shape s, const T &v) : Mat2D(int64_t(s.x), int64_t(s.y), v) {}
  Mat2D(const std::initializer_list<std::initializer_list<T>> &rows) : shape_(0, 0) {
    for (const auto &r : rows) push_back(std::vector<T>(r));
  }

  T &at(int64_t row, int64_t col) noexcept {
    assert(row >= 0 && uint64_t(row) < shape_.y && col >= 0 && uint64_t(col) < shape_.x);
    return data_[size_t(row) * shape_.x + size_t(col)];
  }
  const T &at(int64_t row, int64_t col) const noexcept {
    assert(row >= 0 && uint64_t(row) < shape_.y && col >= 0 && uint64_t(col) < shape_.x);
    return data_[size_t(row) * shape_.x + size_t(col)];
  }
  Row operator[](int64_t row) noexcept { return Row(data_.data() + size_t(row) * shape_.x, int64_t(shape_.x)); }
  ConstRow operator[](int64_t row) const noexcept {
    return ConstRow(data_.data() + size_t(row) * shape_.x, int64_t(shape_.x));
  }

  const Shape &shape() const noexcept { return shape_; }
  T *data() noexcept { return data_.data(); }
  const T *data() const noexcept { return data_.data(); }

  /* grow or shrink to cols x rows, preserving the overlapping top-left block */
  void resize(int64_t cols, int64_t rows) {
    Mat2D next(cols, rows);
    const int64_t cr = std::min<int64_t>(rows, int64_t(shape_.y));
    const int64_t cc = std::min<int64_t>(cols, int64_t(shape_.x));
    for (int64_t i = 0; i < cr; ++i)
      for (int64_t j = 0; j < cc; ++j) next.at(i, j) = at(i, j);
    data_.swap(next.data_);
    shape_ = next.shape_;
  }

  void push_back(const std::vector<T> &row) {
    assert(shape_.y == 0 || row.size() == shape_.x);
    resize(int64_t(row.size()), int64_t(shape_.y) + 1);
    (*this)[int64_t(shape_.y) - 1] = row;
  }

  bool operator==(const Mat2D &r) const noexcept { return shape_ == r.shape_ && data_ == r.data_; }
  bool operator!=(const Mat2D &r) const noexcept { return !(*this == r); }

  template <typename S> Mat2D &operator/=(const S &s) {
    for (auto &e : data_) e /= s;
    return *this;
  }
  template <typename S> Mat2D &operator*=(const S &s) {
    for (auto &e : data_) e *= s;
    return *this;
  }
};

template <typename T> std::ostream &operator<<(std::ostream &os, const Mat2D<T> &m) {
  for (uint64_t i = 0; i < m.shape().y; ++i) {
    for (uint64_t j = 0; j < m.shape().x; ++j) os << (j ? " " : "") << m.at(i, j);
    os << "\n";
  }
  return os;
}

/* element-wise 1/x with 0 -> +inf (bandwidth matrix -> distance matrix) */
inline Mat2D<double> make_reciprocal(const Mat2D<double> &m) {
  Mat2D<double> out(m.shape());
  for (uint64_t i = 0; i < m.shape().y; ++i)
    for (uint64_t j = 0; j < m.shape().x; ++j) {
      const double e = m.at(i, j);
      out.at(i, j) = (e == 0) ? std::numeric_limits<double>::infinity() : 1.0 / e;
    }
  return out;
}

/* out[i][j] = m[map[i]][map[j]] */
template <typename T> Mat2D<T> permute(const Mat2D<T> &m, const std::vector<size_t> &map) {
  assert(m.shape().x == m.shape().y && m.shape().x == map.size());
  Mat2D<T> out(m.shape());
  for (size_t i = 0; i < map.size(); ++i)
    for (size_t j = 0; j < map.size(); ++j) out.at(i, j) = m.at(map[i], map[j]);
  return out;
}
