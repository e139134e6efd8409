#pragma once
/* Thin wrappers over the Blackwell/Hopper asynchronous-copy PTX used by the stencil kernels: mbarrier
   (init / arrive / expect_tx / parity wait) and cp.async.bulk.tensor.3d (TMA tile load into shared memory), plus the
   host-side tensor-map encoder fetched through the runtime so that the library does not link against libcuda. */
#include <cstdlib>

#include <cuda.h>
#include <cuda_runtime.h>

#include "stencil/cuda_runtime.hpp"
#include "stencil/logging.hpp"

namespace stencil {
namespace tma {

#if defined(__CUDACC__)
__device__ __forceinline__ unsigned smem_u32(const void *p) { return unsigned(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
               "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)),
               "r"(parity)
               : "memory");
}
/* one 3-D box: coordinates in elements (x fastest); completion is signalled on `bar` as transaction bytes */
__device__ __forceinline__ void load_3d(void *dst, const CUtensorMap *map, int cx, int cy, int cz, unsigned long long *bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                   smem_u32(dst)),
               "l"(map), "r"(smem_u32(bar)), "r"(cx), "r"(cy), "r"(cz)
               : "memory");
}
#endif

/* Keep a small, hot table (the per-launch tile table: every CTA starts by reading its entry, and nothing can be issued
   before it arrives) resident in L2 across iterations that stream gigabytes through it: persisting access-policy window
   on the stream that runs the kernels. Best effort: failures (e.g. MPS without persisting L2) are ignored. */
inline void keep_in_l2(cudaStream_t stream, const void *base, size_t bytes) {
  if (!base || !bytes || std::getenv("STENCIL_NO_L2_KEEP")) return;
  (void)cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 4u << 20);
  cudaStreamAttrValue v;
  v.accessPolicyWindow.base_ptr = const_cast<void *>(base);
  v.accessPolicyWindow.num_bytes = bytes;
  v.accessPolicyWindow.hitRatio = 1.0f;
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  (void)cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &v);
  (void)cudaGetLastError();
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CUDA_RUNTIME(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (!p || q != cudaDriverEntryPointSuccess) LOG_FATAL("cuTensorMapEncodeTiled is not available from this driver");
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

/* tensor over one padded field buffer: dims (row pitch, rows, planes) in elements, box (bx, by, 1); out-of-range parts
   of a box are zero-filled by the hardware */
inline CUtensorMap make_plane_map(void *bufferStart, CUtensorMapDataType type, size_t elemSize, long long pitchX, long long rows, long long planes,
                                  unsigned boxX, unsigned boxY) {
  CUtensorMap m;
  const cuuint64_t dims[3] = {cuuint64_t(pitchX), cuuint64_t(rows), cuuint64_t(planes)};
  const cuuint64_t strides[2] = {cuuint64_t(pitchX) * elemSize, cuuint64_t(pitchX) * cuuint64_t(rows) * elemSize};
  const cuuint32_t box[3] = {boxX, boxY, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult rc = encode_tiled()(&m, type, 3, bufferStart, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) LOG_FATAL("cuTensorMapEncodeTiled failed with " << int(rc));
  return m;
}

} // namespace tma
} // namespace stencil
