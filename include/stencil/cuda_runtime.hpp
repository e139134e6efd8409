#pragma once
/* CUDA_RUNTIME(stmt): check a CUDA runtime call; on error print "file:line: CUDA error: <string>" and exit(-1).
   Parity: reference include/stencil/cuda_runtime.hpp:8-17. Additionally CUDA_DRIVER() for driver-API calls and a
   non-fatal CUDA_RUNTIME_SOFT() that returns the error (the reference only has the fatal form). */
#include <cstdio>
#include <cstdlib>

#include <cuda_runtime.h>

enum class CudaErrorsFatal { NO, YES };

namespace stencil {
namespace detail {
inline void check_cuda(cudaError_t e, const char *file, int line) {
  if (e != cudaSuccess) {
    std::fprintf(stderr, "%s:%d: CUDA error: %s (%s)\n", file, line, cudaGetErrorString(e), cudaGetErrorName(e));
    std::exit(-1);
  }
}
inline cudaError_t soft_cuda(cudaError_t e, const char *file, int line) {
  if (e != cudaSuccess) {
    std::fprintf(stderr, "%s:%d: CUDA error (non-fatal): %s\n", file, line, cudaGetErrorString(e));
    (void)cudaGetLastError();
  }
  return e;
}
} // namespace detail
} // namespace stencil

#define CUDA_RUNTIME(stmt) ::stencil::detail::check_cuda((stmt), __FILE__, __LINE__);
#define CUDA_RUNTIME_SOFT(stmt) ::stencil::detail::soft_cuda((stmt), __FILE__, __LINE__)
