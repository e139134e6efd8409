/* Micro-benchmark behind the thin-row ("x face") path of box_copy: copy the first 3 cells of every row of a pitched
   fp32 array (512 x 512 rows of 576 floats, several quantities) into the 32-byte sector that starts one row length
   further (halo + padding), with different access strategies. Prints microseconds per pass; run under ncu for bytes.
     nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o xface tools/microbench/xface.cu -lcuda */
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { std::printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); std::exit(1); } } while (0)

constexpr int kPitch = 576, kRowsY = 514, kPlanes = 514, kSx = 512, kLead = 32; // floats; interior x=0 at kLead (128 B aligned)

__device__ __forceinline__ unsigned ld_stream(const unsigned *p) {
  unsigned v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

/* (a) what box_copy does today: one row per thread, three scalar streaming loads, two 16-byte stores */
__global__ void k_scalar(const float *src, float *dst, long long rows) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    const unsigned *a = reinterpret_cast<const unsigned *>(src + r * kPitch + kLead);
    const unsigned w0 = ld_stream(a), w1 = ld_stream(a + 1), w2 = ld_stream(a + 2);
    uint4 *b = reinterpret_cast<uint4 *>(dst + r * kPitch + kLead + kSx);
    b[0] = make_uint4(w0, w1, w2, 0);
    b[1] = make_uint4(0, 0, 0, 0);
  }
}

/* (b) one 16-byte load, one 32-byte store per row */
__global__ void k_vec(const float *src, float *dst, long long rows) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    uint4 v;
    const float *a = src + r * kPitch + kLead;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a));
    float *b = dst + r * kPitch + kLead + kSx;
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(b), "r"(v.x), "r"(v.y), "r"(v.z), "r"(0), "r"(0), "r"(0), "r"(0), "r"(0) : "memory");
  }
}

/* (b2) as (b) with default-cached loads (ld.global) */
__global__ void k_vec_ca(const float *src, float *dst, long long rows) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    const uint4 v = *reinterpret_cast<const uint4 *>(src + r * kPitch + kLead);
    float *b = dst + r * kPitch + kLead + kSx;
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(b), "r"(v.x), "r"(v.y), "r"(v.z), "r"(0), "r"(0), "r"(0), "r"(0), "r"(0) : "memory");
  }
}

/* (c) TMA: a box of 8 floats x R rows into shared memory, rearranged, stored back with a TMA store box */
constexpr int kBoxRows = 128;
__device__ __forceinline__ unsigned smem_u32(const void *p) { return unsigned(__cvta_generic_to_shared(p)); }
__global__ void __launch_bounds__(128) k_tma(const __grid_constant__ CUtensorMap mapSrc, const __grid_constant__ CUtensorMap mapDst, int nboxY, int planes) {
  __shared__ __align__(128) float in[kBoxRows * 8];
  __shared__ __align__(128) float out[kBoxRows * 8];
  __shared__ __align__(8) unsigned long long bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned phase = 0;
  const long long total = (long long)nboxY * planes;
  for (long long t = blockIdx.x; t < total; t += gridDim.x) {
    const int z = int(t / nboxY), y0 = int(t % nboxY) * kBoxRows;
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(kBoxRows * 32) : "memory");
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(in)),
                   "l"(&mapSrc), "r"(smem_u32(&bar)), "r"(kLead), "r"(y0), "r"(z)
                   : "memory");
    }
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(&bar)), "r"(phase) : "memory");
    phase ^= 1u;
    /* row r: payload words 0..2 of the source box go to words 0..2 of the destination sector, the rest is padding */
    const int r = threadIdx.x;
    const float4 v = *reinterpret_cast<const float4 *>(in + r * 8);
    *reinterpret_cast<float4 *>(out + r * 8) = make_float4(v.x, v.y, v.z, 0.f);
    *reinterpret_cast<float4 *>(out + r * 8 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(&mapDst), "r"(smem_u32(out)), "r"(kLead + kSx),
                   "r"(y0), "r"(z)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); // smem may be overwritten
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                             const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

CUtensorMap make_map(void *base, CUtensorMapL2promotion promo) {
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  CUtensorMap m;
  const cuuint64_t dims[3] = {kPitch, kRowsY, kPlanes};
  const cuuint64_t strides[2] = {kPitch * 4ull, kPitch * 4ull * kRowsY};
  const cuuint32_t box[3] = {8, kBoxRows, 1}, es[3] = {1, 1, 1};
  const CUresult rc = reinterpret_cast<EncodeFn>(fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                     CU_TENSOR_MAP_SWIZZLE_NONE, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) { std::printf("encode failed %d\n", int(rc)); std::exit(1); }
  return m;
}

int main(int argc, char **argv) {
  const int nq = argc > 1 ? std::atoi(argv[1]) : 4, reps = 10;
  const size_t elems = size_t(kPitch) * kRowsY * kPlanes;
  std::vector<float *> bufs(static_cast<size_t>(nq));
  for (auto &b : bufs) {
    CK(cudaMalloc(&b, elems * 4 + 256));
    CK(cudaMemset(b, 0, elems * 4 + 256));
  }
  const long long rows = (long long)kRowsY * kPlanes;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  auto timeit = [&](const char *name, auto launch) {
    for (int w = 0; w < 2; ++w)
      for (float *b : bufs) launch(b);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r)
      for (float *b : bufs) launch(b);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    CK(cudaGetLastError());
    std::printf("%-28s %8.2f us per pass over %d quantities (%lld rows each)\n", name, ms * 1e3 / reps, nq, rows);
  };
  timeit("scalar 3x4B ld, 2x16B st", [&](float *b) { k_scalar<<<148 * 8, 256>>>(b, b, rows); });
  timeit("16B nc ld, 32B st", [&](float *b) { k_vec<<<148 * 8, 256>>>(b, b, rows); });
  timeit("16B ld, 32B st", [&](float *b) { k_vec_ca<<<148 * 8, 256>>>(b, b, rows); });
  for (int promo = 0; promo < 3; ++promo) {
    const CUtensorMapL2promotion p = promo == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : (promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
    std::vector<CUtensorMap> maps;
    for (float *b : bufs) maps.push_back(make_map(b, p));
    char name[64];
    std::snprintf(name, sizeof(name), "TMA box ld + TMA st, promo %d", promo);
    size_t i = 0;
    timeit(name, [&](float *b) {
      (void)b;
      const CUtensorMap &m = maps[i++ % maps.size()];
      k_tma<<<148 * 8, 128>>>(m, m, (kRowsY + kBoxRows - 1) / kBoxRows, kPlanes);
    });
  }
  return 0;
}
