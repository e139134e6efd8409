/* Field reductions. See include/stencil/reduce.hpp. */
#include "stencil/reduce.hpp"

#include <algorithm>

#include "stencil/cuda_runtime.hpp"

namespace stencil {

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
    reduce_sum_f32_kernel(const float *__restrict__ src, long long pitchX, long long pitchXY, int ex, int ey, int ez,
                          double *partials, unsigned long long *ticket, double *out) {
  /* rows are distributed round-robin over CTAs; a thread strides along x */
  double acc = 0;
  const long long rows = (long long)ey * ez;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const float *p = src + (r / ey) * pitchXY + (r % ey) * pitchX;
    for (int x = threadIdx.x; x < ex; x += kThreads) acc += double(p[x]);
  }
  /* CTA reduction in a fixed tree order */
  __shared__ double sh[kThreads];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kThreads / 2; s > 0; s >>= 1) {
    if (int(threadIdx.x) < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  __shared__ bool last;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = sh[0];
    __threadfence();
    last = atomicAdd(ticket, 1ull) == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    /* the last CTA adds all partials in index order: same result whatever the arrival order was */
    __threadfence();
    double a = 0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += kThreads) a += partials[i];
    sh[threadIdx.x] = a;
    __syncthreads();
    for (int s = kThreads / 2; s > 0; s >>= 1) {
      if (int(threadIdx.x) < s) sh[threadIdx.x] += sh[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      *out = sh[0];
      *ticket = 0; // ready for the next launch
    }
  }
}

} // namespace

void reduce_sum_f32(const float *src, const Dim3 &pitch, const Dim3 &ext, void *scratch, double *out, cudaStream_t stream) {
  double *partials = static_cast<double *>(scratch);
  unsigned long long *ticket = reinterpret_cast<unsigned long long *>(partials + kReduceMaxBlocks);
  const long long rows = std::max<long long>(1, ext.y * ext.z);
  const unsigned blocks = unsigned(std::min<long long>(rows, kReduceMaxBlocks));
  reduce_sum_f32_kernel<<<blocks, kThreads, 0, stream>>>(src, pitch.x, pitch.x * pitch.y, int(ext.x), int(ext.y), int(ext.z), partials,
                                                         ticket, out);
  CUDA_RUNTIME(cudaGetLastError());
}

double reduce_sum_f32_sync(const float *src, const Dim3 &pitch, const Dim3 &ext, cudaStream_t stream) {
  char *buf = nullptr;
  CUDA_RUNTIME(cudaMalloc(&buf, kReduceScratchBytes + sizeof(double)));
  CUDA_RUNTIME(cudaMemsetAsync(buf, 0, kReduceScratchBytes + sizeof(double), stream));
  double *out = reinterpret_cast<double *>(buf + kReduceScratchBytes);
  reduce_sum_f32(src, pitch, ext, buf, out, stream);
  double h = 0;
  CUDA_RUNTIME(cudaMemcpyAsync(&h, out, sizeof(double), cudaMemcpyDeviceToHost, stream));
  CUDA_RUNTIME(cudaStreamSynchronize(stream));
  CUDA_RUNTIME(cudaFree(buf));
  return h;
}

} // namespace stencil
