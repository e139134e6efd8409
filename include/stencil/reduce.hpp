#pragma once
/* Deterministic field reductions (not in the reference, which only dumps fields to CSV): the residual / checksum
   a solver loop reads back each step. Two-stage: every CTA writes one fp64 partial to a scratch array, the last
   CTA to finish (ticket counter) adds the partials in index order, so the result does not depend on scheduling. */
#include <cstddef>

#include <cuda_runtime.h>

#include "stencil/dim3.hpp"

namespace stencil {

/* upper bound on CTAs (= scratch partials) a reduction launch uses */
constexpr unsigned kReduceMaxBlocks = 1184; // 148 SMs x 8 CTAs of 256 threads
/* bytes of scratch a reduction needs: partials + ticket counter */
constexpr size_t kReduceScratchBytes = (kReduceMaxBlocks + 1) * sizeof(double);

/* Sum of a box of an fp32 pitched array, accumulated in fp64.
   src      device pointer to element (0,0,0) of the box
   pitch    elements: x = row pitch, y = rows per plane
   ext      box extent in elements
   scratch  device memory of kReduceScratchBytes; its last 8 bytes (the ticket) must be zero at launch and are reset
            to zero by the kernel, so one cleared buffer serves any number of consecutive launches on one stream
   out      device pointer receiving the sum
   Enqueues on `stream`; no allocation, no synchronisation. */
void reduce_sum_f32(const float *src, const Dim3 &pitch, const Dim3 &ext, void *scratch, double *out, cudaStream_t stream);

/* convenience: allocates scratch + result, runs, synchronises, frees; returns the sum */
double reduce_sum_f32_sync(const float *src, const Dim3 &pitch, const Dim3 &ext, cudaStream_t stream = 0);

} // namespace stencil
