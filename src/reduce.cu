/* Field reductions. See include/stencil/reduce.hpp. */
#include "stencil/reduce.hpp"

#include <algorithm>
#include <climits>
#include <cstdio>
#include <mutex>

#include "stencil_c.h"

#include "stencil/cuda_runtime.hpp"

namespace stencil {

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
    reduce_sum_f32_kernel(const float *__restrict__ src, long long pitchX, long long pitchXY, long long ex, long long ey, long long ez,
                          double *partials, unsigned long long *ticket, double *out) {
  /* rows are distributed round-robin over CTAs; a thread strides along x (64-bit indices: a row may be INT_MAX cells) */
  double acc = 0;
  const long long rows = ey * ez;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const float *p = src + (r / ey) * pitchXY + (r % ey) * pitchX;
    for (long long x = threadIdx.x; x < ex; x += kThreads) acc += double(p[x]);
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
  reduce_sum_f32_kernel<<<blocks, kThreads, 0, stream>>>(src, pitch.x, pitch.x * pitch.y, (long long)ext.x, (long long)ext.y, (long long)ext.z,
                                                         partials, ticket, out);
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

/* ---------------------------------------------------------------------------------------------------------------
   Allocation-free variant of the field sum.

   Who needs it: code that may not call cudaMalloc/cudaFree at the point where it wants the sum. In this library
   that is, first of all, a monitor inside a free-running fused iteration loop: cudaMalloc and cudaFree synchronise
   the whole device, which would stall the per-GPU kernels that are waiting on each other's NVLink flags; second,
   solvers that size their subdomains to fill HBM and have no room left for per-call temporaries; third, callers
   that run under an allocator they do not control (another framework's memory pool).

   So the two-stage reduction's scratch — one fp64 partial per CTA plus the ticket counter — does not come from an
   allocator here: it is the file-scope device array below, present once per GPU that loads this library and held
   until the process exits (STENCIL_NOALLOC_SCRATCH_BYTES = 9728 bytes).

   In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel can use, so the device
   address is looked up with cudaGetSymbolAddress — once per device, with that device current — and cached. */
static __device__ __align__(256) unsigned char g_noalloc_scratch[STENCIL_NOALLOC_SCRATCH_BYTES];

static_assert(STENCIL_NOALLOC_SCRATCH_BYTES >= stencil::kReduceScratchBytes, "scratch array too small for the reduction");

namespace {
constexpr int kMaxDevices = 64;
void *g_noallocAddr[kMaxDevices] = {nullptr};
std::mutex g_noallocMu;

/* device address of g_noalloc_scratch on the CURRENT device (first call per device resolves and caches it) */
int noalloc_scratch(void **out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    std::fprintf(stderr, "stencil_reduce_sum_f32_noalloc: cannot identify the current device (%s)\n", cudaGetErrorString(e));
    return -10;
  }
  std::lock_guard<std::mutex> lk(g_noallocMu);
  if (!g_noallocAddr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      std::fprintf(stderr, "stencil_reduce_sum_f32_noalloc: cudaGetSymbolAddress failed on device %d (%s)\n", dev, cudaGetErrorString(e));
      return -11;
    }
    g_noallocAddr[dev] = p;
  }
  *out = g_noallocAddr[dev];
  return 0;
}
} // namespace

extern "C" int stencil_reduce_sum_f32_noalloc(const float *src, const int64_t pitch[3], const int64_t ext[3], void *stream, double *sum_out) {
  if (!src || !sum_out || !pitch || !ext) {
    std::fprintf(stderr, "stencil_reduce_sum_f32_noalloc: null argument\n");
    return -1;
  }
  if (ext[0] < 0 || ext[1] < 0 || ext[2] < 0 || pitch[0] < ext[0] || pitch[1] < ext[1]) {
    std::fprintf(stderr, "stencil_reduce_sum_f32_noalloc: negative extent or pitch smaller than the box\n");
    return -2;
  }
  /* the stated maximum of this entry point (the kernel itself indexes with 64 bits): reject anything beyond, loudly */
  if (ext[0] > STENCIL_NOALLOC_MAX_ROW_CELLS || (ext[1] > 0 && ext[2] > STENCIL_NOALLOC_MAX_ROWS / ext[1])) {
    std::fprintf(stderr,
                 "stencil_reduce_sum_f32_noalloc: box %lld x %lld x %lld exceeds the supported maximum (%lld cells per row, %lld rows);"
                 " use stencil_reduce_sum_f32, which allocates\n",
                 (long long)ext[0], (long long)ext[1], (long long)ext[2], (long long)STENCIL_NOALLOC_MAX_ROW_CELLS,
                 (long long)STENCIL_NOALLOC_MAX_ROWS);
    return -3;
  }
  void *scratch = nullptr;
  if (const int rc = noalloc_scratch(&scratch)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  /* the array keeps whatever the previous call left behind; the partials are overwritten before they are read, but
     the ticket must start at zero (a call that was aborted, or one on another stream, may have left it anywhere) */
  unsigned char *ticket = static_cast<unsigned char *>(scratch) + stencil::kReduceMaxBlocks * sizeof(double);
  cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned long long), st);
  if (e != cudaSuccess) {
    std::fprintf(stderr, "stencil_reduce_sum_f32_noalloc: clearing the ticket failed (%s)\n", cudaGetErrorString(e));
    return -12;
  }
  stencil::reduce_sum_f32(src, Dim3(pitch[0], pitch[1], pitch[2]), Dim3(ext[0], ext[1], ext[2]), scratch, sum_out, st);
  return 0;
}
