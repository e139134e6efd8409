#pragma once
/* unpack / translate / multi_* box copies.
   Parity: reference include/stencil/copy.cuh:13-146 (K4 multi_translate, K5 translate, K6 unpack, K7 multi_pack /
   multi_unpack). Same naming and argument order, but these are host functions that enqueue the sm_100a mover
   (src/box_copy.cu) instead of __global__ symbols with caller-made grids; pointer/size tables are HOST arrays
   (descriptors travel as kernel parameters, no device-side tables to keep in sync). Pitches are in elements. */
#include "stencil/copy_plan.hpp"
#include "stencil/pack_kernel.cuh"

namespace stencil {

/* dense src -> box of a pitched dst */
inline void unpack(void *dst, const Dim3 &dstSize, const Dim3 &dstPos, const Dim3 &dstExtent, const void *src,
                   size_t elemSize, cudaStream_t stream = 0) {
  if (dstExtent.flatten() <= 0) return;
  SegBatch b;
  b.n = 1;
  b.seg[0] = make_seg(BoxView::dense(const_cast<void *>(src), dstExtent, elemSize), Dim3(0, 0, 0),
                      BoxView::pitched(dst, dst, dstSize, elemSize), dstPos, dstExtent);
  launch_box_copy(b, 0, stream);
}

/* box of a pitched src -> box of a pitched dst */
inline void translate(void *dst, const Dim3 &dstPos, const Dim3 &dstSize, const void *src, const Dim3 &srcPos,
                      const Dim3 &srcSize, const Dim3 &extent, size_t elemSize, cudaStream_t stream = 0) {
  if (extent.flatten() <= 0) return;
  SegBatch b;
  b.n = 1;
  void *s = const_cast<void *>(src);
  b.seg[0] = make_seg(BoxView::pitched(s, s, srcSize, elemSize), srcPos, BoxView::pitched(dst, dst, dstSize, elemSize),
                      dstPos, extent);
  launch_box_copy(b, 0, stream);
}

/* the same box of n quantities (one launch per SegBatch::kMax quantities) */
inline void multi_translate(void *const *dsts, const Dim3 &dstPos, const Dim3 &dstSize, void *const *srcs,
                            const Dim3 &srcPos, const Dim3 &srcSize, const Dim3 &extent, const size_t *elemSizes, size_t n,
                            cudaStream_t stream = 0) {
  if (extent.flatten() <= 0) return;
  for (size_t base = 0; base < n; base += SegBatch::kMax) {
    SegBatch b;
    b.n = unsigned(n - base < size_t(SegBatch::kMax) ? n - base : size_t(SegBatch::kMax));
    for (unsigned i = 0; i < b.n; ++i) {
      const size_t q = base + i;
      /* per-quantity pitches scale with nothing: sizes are in elements, so one Dim3 serves every quantity */
      b.seg[i] = make_seg(BoxView::pitched(srcs[q], srcs[q], srcSize, elemSizes[q]), srcPos,
                          BoxView::pitched(dsts[q], dsts[q], dstSize, elemSizes[q]), dstPos, extent);
    }
    launch_box_copy(b, 0, stream);
  }
}

/* pack the same box of n quantities back to back into dst at the given byte offsets */
inline void multi_pack(void *dst, const size_t *offsets, void *const *srcs, const Dim3 &srcSize, const Dim3 &srcPos,
                       const Dim3 &srcExtent, const size_t *elemSizes, size_t n, cudaStream_t stream = 0) {
  if (srcExtent.flatten() <= 0) return;
  for (size_t base = 0; base < n; base += SegBatch::kMax) {
    SegBatch b;
    b.n = unsigned(n - base < size_t(SegBatch::kMax) ? n - base : size_t(SegBatch::kMax));
    for (unsigned i = 0; i < b.n; ++i) {
      const size_t q = base + i;
      b.seg[i] = make_seg(BoxView::pitched(srcs[q], srcs[q], srcSize, elemSizes[q]), srcPos,
                          BoxView::dense(static_cast<char *>(dst) + offsets[q], srcExtent, elemSizes[q]), Dim3(0, 0, 0),
                          srcExtent);
    }
    launch_box_copy(b, 0, stream);
  }
}

inline void multi_unpack(void *const *dsts, const Dim3 &dstSize, const Dim3 &dstPos, const Dim3 &dstExtent,
                         const void *src, const size_t *offsets, const size_t *elemSizes, size_t n,
                         cudaStream_t stream = 0) {
  if (dstExtent.flatten() <= 0) return;
  for (size_t base = 0; base < n; base += SegBatch::kMax) {
    SegBatch b;
    b.n = unsigned(n - base < size_t(SegBatch::kMax) ? n - base : size_t(SegBatch::kMax));
    for (unsigned i = 0; i < b.n; ++i) {
      const size_t q = base + i;
      char *from = const_cast<char *>(static_cast<const char *>(src)) + offsets[q];
      b.seg[i] = make_seg(BoxView::dense(from, dstExtent, elemSizes[q]), Dim3(0, 0, 0),
                          BoxView::pitched(dsts[q], dsts[q], dstSize, elemSizes[q]), dstPos, dstExtent);
    }
    launch_box_copy(b, 0, stream);
  }
}

} // namespace stencil
