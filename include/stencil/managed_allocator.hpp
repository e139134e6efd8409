#pragma once
/* ManagedAllocator<T>: std-allocator-shaped wrapper over cudaMallocManaged (usable with std::vector).
   Parity: reference include/stencil/managed_allocator.hpp:7-64. */
#include <cstddef>
#include <new>

#include "stencil/cuda_runtime.hpp"

template <class T> class ManagedAllocator {
  int dev_;

public:
  typedef T value_type;
  explicit ManagedAllocator(int dev = 0) noexcept : dev_(dev) {}
  template <class U> ManagedAllocator(const ManagedAllocator<U> &o) noexcept : dev_(o.device()) {}
  int device() const noexcept { return dev_; }

  T *allocate(std::size_t n) {
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(dev_));
    void *p = nullptr;
    const cudaError_t e = cudaMallocManaged(&p, n * sizeof(T));
    CUDA_RUNTIME(cudaSetDevice(prev));
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      throw std::bad_alloc();
    }
    return static_cast<T *>(p);
  }
  /* largest n allocate(n) could be asked for (std allocator requirement kept by the reference, :38) */
  std::size_t max_size() const noexcept { return static_cast<std::size_t>(-1) / sizeof(T); }
  void deallocate(T *p, std::size_t) noexcept {
    if (p) cudaFree(p);
  }
  template <class U> bool operator==(const ManagedAllocator<U> &o) const noexcept { return dev_ == o.device(); }
  template <class U> bool operator!=(const ManagedAllocator<U> &o) const noexcept { return dev_ != o.device(); }
};
