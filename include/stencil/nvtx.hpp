#pragma once
/* Named ranges for timeline tools. The reference brackets its setup phases, every transport's send / recv / wait,
   swap() and write_paraview() with NVTX v1 push/pop (stencil.hpp:240-246, src/stencil.cu:35-44,
   :672-861, tx_cuda.cuh:70 ff., src/local_domain.cu:42). Here the same range names are emitted through the
   header-only NVTX 3 (no library to link; a no-op unless a tool injects itself), as a scope guard so that early
   returns cannot leave a range open. -DSTENCIL_NO_NVTX compiles them out. */
#ifndef STENCIL_NO_NVTX
#include <nvtx3/nvToolsExt.h>
#endif

namespace stencil {
namespace nvtx {

class Range {
public:
  explicit Range(const char *name) {
#ifndef STENCIL_NO_NVTX
    nvtxRangePushA(name);
#else
    (void)name;
#endif
  }
  ~Range() {
#ifndef STENCIL_NO_NVTX
    nvtxRangePop();
#endif
  }
  Range(const Range &) = delete;
  Range &operator=(const Range &) = delete;
};

inline void mark(const char *name) {
#ifndef STENCIL_NO_NVTX
  nvtxMarkA(name);
#else
  (void)name;
#endif
}

} // namespace nvtx
} // namespace stencil

#define STENCIL_NVTX_CAT2(a, b) a##b
#define STENCIL_NVTX_CAT(a, b) STENCIL_NVTX_CAT2(a, b)
#define STENCIL_RANGE(name) ::stencil::nvtx::Range STENCIL_NVTX_CAT(stencilRange_, __LINE__)(name)
