/* RcStream implementation. Reference: src/rcstream.cpp:5-46. */
#include "stencil/rcstream.hpp"

#include "stencil/cuda_runtime.hpp"

namespace {
/* keep the caller's current device unchanged around stream creation/destruction */
struct DeviceGuard {
  int prev = 0;
  explicit DeviceGuard(int dev) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != dev) CUDA_RUNTIME(cudaSetDevice(dev));
  }
  ~DeviceGuard() { cudaSetDevice(prev); }
};
} // namespace

RcStream::RcStream(int dev, Priority p) {
  DeviceGuard g(dev);
  cudaStream_t s = nullptr;
  if (p == Priority::HIGH) {
    int least = 0, greatest = 0; // numerically lowest value = highest priority
    CUDA_RUNTIME(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CUDA_RUNTIME(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, greatest));
  } else {
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  }
  h_ = std::shared_ptr<Handle>(new Handle{s, dev}, [](Handle *h) {
    int prev = 0;
    if (cudaGetDevice(&prev) == cudaSuccess) {
      cudaSetDevice(h->dev);
      cudaStreamDestroy(h->stream);
      cudaSetDevice(prev);
    }
    delete h;
  });
}

void RcStream::synchronize() const { CUDA_RUNTIME(cudaStreamSynchronize(get())); }
