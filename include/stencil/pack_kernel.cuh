#pragma once
/* pack: copy a 3-D box of a pitched array into a dense buffer.
   Parity: reference include/stencil/pack_kernel.cuh:5-46 (`pack_kernel`, K1). The reference exposes a __global__
   the caller launches with a hand-made grid; here it is a host function that enqueues the vectorised sm_100a mover
   (src/box_copy.cu) on `stream` — grid shaping is the library's job. `srcSize` is the source pitch in elements
   (x = row pitch, y = rows per plane), exactly what LocalDomain::pitch(qi) returns. */
#include "stencil/copy_plan.hpp"

namespace stencil {

inline void pack(void *dst, const void *src, const Dim3 &srcSize, const Dim3 &srcPos, const Dim3 &srcExtent,
                 size_t elemSize, cudaStream_t stream = 0) {
  if (srcExtent.flatten() <= 0) return;
  SegBatch b;
  b.n = 1;
  void *s = const_cast<void *>(src);
  b.seg[0] = make_seg(BoxView::pitched(s, s, srcSize, elemSize), srcPos, BoxView::dense(dst, srcExtent, elemSize),
                      Dim3(0, 0, 0), srcExtent);
  launch_box_copy(b, 0, stream);
}

} // namespace stencil

/* reference spelling */
inline void pack_kernel(void *dst, const void *src, const Dim3 &srcSize, const Dim3 &srcPos, const Dim3 &srcExtent,
                        size_t elemSize, cudaStream_t stream = 0) {
  stencil::pack(dst, src, srcSize, srcPos, srcExtent, elemSize, stream);
}
