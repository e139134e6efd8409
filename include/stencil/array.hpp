#pragma once
/* Array<T, device>: minimal owning array (resize/size/data/[]/swap).
   Parity: reference include/stencil/array.hpp:3-84. There the GPU specialisation is dead code (guarded by an
   undefined macro) and Array<T,gpu> silently lives on the host; here device_type::gpu really allocates device
   memory when compiled with STENCIL_USE_CUDA (the default). */
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <utility>

#ifndef STENCIL_USE_CUDA
#define STENCIL_USE_CUDA 1
#endif
#if STENCIL_USE_CUDA == 1
#include "stencil/cuda_runtime.hpp"
#endif

enum class device_type { cpu, gpu };

template <typename T, device_type D = device_type::cpu> class Array {
  T *data_ = nullptr;
  size_t size_ = 0;

public:
  Array() {}
  explicit Array(size_t n) { resize(n); }
  Array(size_t n, const T &v) {
    resize(n);
    for (size_t i = 0; i < n; ++i) data_[i] = v;
  }
  Array(const Array &o) {
    resize(o.size_);
    if (size_) std::memcpy(data_, o.data_, size_ * sizeof(T));
  }
  Array(Array &&o) noexcept { swap(o); }
  Array &operator=(Array o) noexcept {
    swap(o);
    return *this;
  }
  ~Array() { std::free(data_); }
  void swap(Array &o) noexcept {
    std::swap(data_, o.data_);
    std::swap(size_, o.size_);
  }
  /* contents are preserved up to min(old, new) elements */
  void resize(size_t n) {
    if (n == size_) return;
    T *p = n ? static_cast<T *>(std::malloc(n * sizeof(T))) : nullptr;
    if (p && data_) std::memcpy(p, data_, (n < size_ ? n : size_) * sizeof(T));
    std::free(data_);
    data_ = p;
    size_ = n;
  }
  size_t size() const noexcept { return size_; }
  T *data() noexcept { return data_; }
  const T *data() const noexcept { return data_; }
  T &operator[](size_t i) noexcept { return data_[i]; }
  const T &operator[](size_t i) const noexcept { return data_[i]; }
};

#if STENCIL_USE_CUDA == 1
/* device-resident storage on the current device; element access is by explicit copies */
template <typename T> class Array<T, device_type::gpu> {
  T *data_ = nullptr;
  size_t size_ = 0;

public:
  Array() {}
  explicit Array(size_t n) { resize(n); }
  Array(const Array &) = delete;
  Array &operator=(const Array &) = delete;
  Array(Array &&o) noexcept { swap(o); }
  Array &operator=(Array &&o) noexcept {
    swap(o);
    return *this;
  }
  ~Array() {
    if (data_) cudaFree(data_);
  }
  void swap(Array &o) noexcept {
    std::swap(data_, o.data_);
    std::swap(size_, o.size_);
  }
  void resize(size_t n) {
    if (n == size_) return;
    T *p = nullptr;
    if (n) CUDA_RUNTIME(cudaMalloc(&p, n * sizeof(T)));
    if (p && data_) CUDA_RUNTIME(cudaMemcpy(p, data_, (n < size_ ? n : size_) * sizeof(T), cudaMemcpyDeviceToDevice));
    if (data_) CUDA_RUNTIME(cudaFree(data_));
    data_ = p;
    size_ = n;
  }
  size_t size() const noexcept { return size_; }
  T *data() noexcept { return data_; }
  const T *data() const noexcept { return data_; }
  void from_host(const T *src, size_t n) { CUDA_RUNTIME(cudaMemcpy(data_, src, n * sizeof(T), cudaMemcpyHostToDevice)); }
  void to_host(T *dst, size_t n) const { CUDA_RUNTIME(cudaMemcpy(dst, data_, n * sizeof(T), cudaMemcpyDeviceToHost)); }
};
#endif
