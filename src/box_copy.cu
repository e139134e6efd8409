/* box_copy_kernel: the sm_100a data mover behind pack / unpack / translate / halo push (kernels N1, N6, N7 of
   SURVEY section 8; replaces reference kernels K1-K7: pack_kernel.cuh:5-46, copy.cuh:13-146, packer.cuh:52-69,194-250).

   Differences from the reference kernels that matter on B200:
   * one launch for ANY number of boxes/quantities/messages (the reference launches one kernel per message and loops
     over quantities serially inside each thread);
   * access width chosen per segment from the actual alignment (16 B vectors for y/z faces whose rows start on the
     128 B-aligned interior; scalar for the 1-3 element rows of x faces), not a per-element switch on elemSize;
   * flat (row, unit) index space per chunk, so thin rows keep all 32 lanes busy (the reference maps a 3-wide face
     onto a 4-wide block and idles 25 % of the lanes; a 1-wide face gets one sector per thread either way);
   * loads are streaming (ld.global.nc.L1::no_allocate) and 4 deep per thread before the stores, so one thread has
     up to 64 B in flight towards NVLink/HBM;
   * 64-bit byte offsets (the reference's pack indices overflow past 4 Gi elements);
   * optional release-signal epilogue so the same launch that stores into a peer's halo also tells it "done". */
#include "stencil/copy_plan.hpp"

#include <algorithm>
#include <cstdio>

#include "stencil/cuda_runtime.hpp"

namespace stencil {

namespace {

template <int BYTES> struct VecT;
template <> struct VecT<16> { typedef uint4 type; };
template <> struct VecT<8> { typedef uint2 type; };
template <> struct VecT<4> { typedef unsigned type; };
template <> struct VecT<2> { typedef unsigned short type; };
template <> struct VecT<1> { typedef unsigned char type; };

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 ld_stream(const uint2 *p) {
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_stream(const unsigned *p) {
  unsigned v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned short ld_stream(const unsigned short *p) { return __ldg(p); }
__device__ __forceinline__ unsigned char ld_stream(const unsigned char *p) { return __ldg(p); }

/* Flag stores are relaxed and follow ONE system-scope fence: a st.release.sys per flag is a fence each, and 26 of
   them in a row by one thread put a ~15 us floor under every exchange. */
__device__ __forceinline__ void st_relaxed_sys(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr int kThreads = 256;

/* A thin row of nw <= 4 words that starts `so` words into an aligned 16-byte window which may be read whole
   (CopySeg::srcWin): the payload arrives with one or two 16-byte requests instead of one per word. */
__device__ __forceinline__ void load_row_window(const char *__restrict__ a, unsigned so, unsigned nw, unsigned (&p)[4]) {
  const uint4 *q = reinterpret_cast<const uint4 *>(a - 4 * (long long)so);
  const uint4 q0 = ld_stream(q);
  uint4 q1 = make_uint4(0u, 0u, 0u, 0u);
  if (so + nw > 4) q1 = ld_stream(q + 1);
  switch (so) { // uniform per segment
  case 0: p[0] = q0.x, p[1] = q0.y, p[2] = q0.z, p[3] = q0.w; break;
  case 1: p[0] = q0.y, p[1] = q0.z, p[2] = q0.w, p[3] = q1.x; break;
  case 2: p[0] = q0.z, p[1] = q0.w, p[2] = q1.x, p[3] = q1.y; break;
  default: p[0] = q0.w, p[1] = q1.x, p[2] = q1.y, p[3] = q1.z; break;
  }
}

/* 32 bytes in one request (sm_100: 256-bit global stores) */
__device__ __forceinline__ void st_sector(void *p, unsigned a, unsigned b, unsigned c, unsigned d, unsigned e, unsigned f, unsigned g, unsigned h) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}

/* one thin row in sector mode: gather the payload words into their place inside 1-2 whole sectors, zero elsewhere */
__device__ __forceinline__ void copy_row_sectors(const CopySeg &s, const char *__restrict__ a, char *__restrict__ b) {
  const unsigned lo = s.padLoWords, nw = s.rowBytes / 4, total = s.totalWords;
  const unsigned *src = reinterpret_cast<const unsigned *>(a);
  unsigned w[16];
  if (s.srcWin) {
    unsigned p[4];
    load_row_window(a, s.srcWin - 1, nw, p);
#pragma unroll
    for (unsigned j = 0; j < 16; ++j) {
      const unsigned d = j - lo; // wraps to a huge value below lo
      w[j] = d < nw ? (d == 0 ? p[0] : (d == 1 ? p[1] : (d == 2 ? p[2] : p[3]))) : 0u;
    }
  } else {
#pragma unroll
    for (unsigned j = 0; j < 16; ++j) w[j] = (j >= lo && j < lo + nw) ? ld_stream(src + (j - lo)) : 0u;
  }
  char *dst = b - 4 * (long long)lo; // 32-byte aligned by construction of sector mode
  st_sector(dst, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
  if (total > 8) st_sector(dst + 32, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]);
}

/* Copy `nrows` consecutive rows of one plane with BYTES-wide accesses. sp/dp point at the first row. No integer
   division anywhere: thin rows (a handful of units, the x faces) are one row per thread; wide rows are one row per
   warp with the lanes striding along the row, two rows in flight per warp. */
template <int BYTES>
__device__ __forceinline__ void copy_rows(const CopySeg &s, const char *__restrict__ sp, char *__restrict__ dp, unsigned nrows) {
  typedef typename VecT<BYTES>::type V;
  const unsigned upr = s.rowBytes / BYTES; // units per row
  if (s.totalWords) {
    for (unsigned r = threadIdx.x; r < nrows; r += kThreads) copy_row_sectors(s, sp + (long long)r * s.srcPitchY, dp + (long long)r * s.dstPitchY);
  } else if (upr <= 8) {
    for (unsigned r = threadIdx.x; r < nrows; r += kThreads) {
      const V *a = reinterpret_cast<const V *>(sp + (long long)r * s.srcPitchY);
      V *b = reinterpret_cast<V *>(dp + (long long)r * s.dstPitchY);
      V v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { // every v[u] is defined on every path: an undefined one stays live across the whole kernel
        v[u] = V();
        if (u < int(upr)) v[u] = ld_stream(a + u);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (u < int(upr)) b[u] = v[u];
    }
  } else {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = kThreads / 32;
    for (unsigned r = warp; r < nrows; r += 2 * nwarps) {
      const unsigned r2 = r + nwarps;
      const V *a0 = reinterpret_cast<const V *>(sp + (long long)r * s.srcPitchY);
      V *b0 = reinterpret_cast<V *>(dp + (long long)r * s.dstPitchY);
      const V *a1 = reinterpret_cast<const V *>(sp + (long long)r2 * s.srcPitchY);
      V *b1 = reinterpret_cast<V *>(dp + (long long)r2 * s.dstPitchY);
      const bool two = r2 < nrows;
      for (unsigned u = lane; u < upr; u += 64) {
        const unsigned u2 = u + 32;
        V v00 = ld_stream(a0 + u), v01 = V(), v10 = V(), v11 = V();
        if (u2 < upr) v01 = ld_stream(a0 + u2);
        if (two) {
          v10 = ld_stream(a1 + u);
          if (u2 < upr) v11 = ld_stream(a1 + u2);
        }
        b0[u] = v00;
        if (u2 < upr) b0[u2] = v01;
        if (two) {
          b1[u] = v10;
          if (u2 < upr) b1[u2] = v11;
        }
      }
    }
  }
}

/* Several whole planes of a box whose planes hold only a few rows (edges and corners): rows are enumerated across
   planes so that all threads stay busy; one division per ROW (never per element). */
template <int BYTES>
__device__ __forceinline__ void copy_planes(const CopySeg &s, const char *__restrict__ sp, char *__restrict__ dp, unsigned nplanes) {
  typedef typename VecT<BYTES>::type V;
  const unsigned upr = s.rowBytes / BYTES, ny = s.ny, total = nplanes * ny;
  if (s.totalWords) {
    for (unsigned p = threadIdx.x; p < total; p += kThreads) {
      const unsigned zi = p / ny, y = p - zi * ny;
      copy_row_sectors(s, sp + (long long)zi * s.srcPitchZ + (long long)y * s.srcPitchY, dp + (long long)zi * s.dstPitchZ + (long long)y * s.dstPitchY);
    }
  } else if (upr <= 8) {
    for (unsigned p = threadIdx.x; p < total; p += kThreads) {
      const unsigned zi = p / ny, y = p - zi * ny;
      const V *a = reinterpret_cast<const V *>(sp + (long long)zi * s.srcPitchZ + (long long)y * s.srcPitchY);
      V *b = reinterpret_cast<V *>(dp + (long long)zi * s.dstPitchZ + (long long)y * s.dstPitchY);
      V v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { // every v[u] is defined on every path: an undefined one stays live across the whole kernel
        v[u] = V();
        if (u < int(upr)) v[u] = ld_stream(a + u);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (u < int(upr)) b[u] = v[u];
    }
  } else {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = kThreads / 32;
    for (unsigned p = warp; p < total; p += nwarps) {
      const unsigned zi = p / ny, y = p - zi * ny;
      const V *a = reinterpret_cast<const V *>(sp + (long long)zi * s.srcPitchZ + (long long)y * s.srcPitchY);
      V *b = reinterpret_cast<V *>(dp + (long long)zi * s.dstPitchZ + (long long)y * s.dstPitchY);
      for (unsigned u = lane; u < upr; u += 32) b[u] = ld_stream(a + u);
    }
  }
}

constexpr unsigned kStageWords = 6144; // 24 KB of staging per CTA

/* `rows` consecutive thin rows (row index -> source address through `rowSrc`) into a contiguous destination */
template <typename RowSrc>
__device__ __forceinline__ void gather_dense(const CopySeg &s, RowSrc rowSrc, char *__restrict__ dp, unsigned rows, unsigned *stage) {
  const unsigned nw = s.rowBytes / 4;
  const unsigned rowsPerPass = kStageWords / nw;
  for (unsigned r0 = 0; r0 < rows; r0 += rowsPerPass) {
    const unsigned n = min(rowsPerPass, rows - r0);
    for (unsigned r = threadIdx.x; r < n; r += kThreads) {
      const unsigned *a = reinterpret_cast<const unsigned *>(rowSrc(r0 + r));
      if (s.srcWin) {
        unsigned p[4];
        load_row_window(reinterpret_cast<const char *>(a), s.srcWin - 1, nw, p);
#pragma unroll
        for (unsigned w = 0; w < 4; ++w)
          if (w < nw) stage[r * nw + w] = p[w];
      } else {
#pragma unroll
        for (unsigned w = 0; w < 16; ++w)
          if (w < nw) stage[r * nw + w] = ld_stream(a + w);
      }
    }
    __syncthreads();
    const unsigned words = n * nw;
    unsigned *out = reinterpret_cast<unsigned *>(dp + (long long)r0 * s.rowBytes);
    if ((reinterpret_cast<unsigned long long>(out) & 15ull) == 0) {
      const unsigned quads = words / 4;
      for (unsigned i = threadIdx.x; i < quads; i += kThreads) reinterpret_cast<uint4 *>(out)[i] = reinterpret_cast<const uint4 *>(stage)[i];
      for (unsigned i = quads * 4 + threadIdx.x; i < words; i += kThreads) out[i] = stage[i];
    } else {
      for (unsigned i = threadIdx.x; i < words; i += kThreads) out[i] = stage[i];
    }
    __syncthreads();
  }
}

template <bool SIGNAL>
__global__ void __launch_bounds__(kThreads, 4)
    box_copy_kernel(const CopySeg *__restrict__ segs, const CopyChunk *__restrict__ chunks, unsigned nchunks, int parity,
                    unsigned *done, SignalSet sig, unsigned epoch) {
  __shared__ CopySeg s;
  __shared__ __align__(16) unsigned stage[kStageWords];
  unsigned cached = 0xffffffffu;
  for (unsigned c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const CopyChunk ch = chunks[c];
    if (ch.seg != cached) { // uniform per CTA
      __syncthreads();
      if (threadIdx.x < sizeof(CopySeg) / sizeof(unsigned))
        reinterpret_cast<unsigned *>(&s)[threadIdx.x] = reinterpret_cast<const unsigned *>(&segs[ch.seg])[threadIdx.x];
      __syncthreads();
      cached = ch.seg;
    }
    const char *src = s.src[parity] + (long long)ch.z * s.srcPitchZ + (long long)ch.y0 * s.srcPitchY;
    char *dst = s.dst[parity] + (long long)ch.z * s.dstPitchZ + (long long)ch.y0 * s.dstPitchY;
    if (s.gatherDense) {
      const long long spy = s.srcPitchY, spz = s.srcPitchZ;
      if (ch.nplanes == 1) {
        gather_dense(s, [=](unsigned r) { return src + (long long)r * spy; }, dst, ch.nrows, stage);
      } else {
        const unsigned ny = s.ny;
        gather_dense(s, [=](unsigned p) {
          const unsigned zi = p / ny, y = p - zi * ny;
          return src + (long long)zi * spz + (long long)y * spy;
        }, dst, unsigned(ch.nplanes) * ny, stage);
      }
    } else if (ch.nplanes == 1) {
      switch (s.vecLog2) {
      case 4: copy_rows<16>(s, src, dst, ch.nrows); break;
      case 3: copy_rows<8>(s, src, dst, ch.nrows); break;
      case 2: copy_rows<4>(s, src, dst, ch.nrows); break;
      case 1: copy_rows<2>(s, src, dst, ch.nrows); break;
      default: copy_rows<1>(s, src, dst, ch.nrows); break;
      }
    } else {
      switch (s.vecLog2) {
      case 4: copy_planes<16>(s, src, dst, ch.nplanes); break;
      case 3: copy_planes<8>(s, src, dst, ch.nplanes); break;
      case 2: copy_planes<4>(s, src, dst, ch.nplanes); break;
      case 1: copy_planes<2>(s, src, dst, ch.nplanes); break;
      default: copy_planes<1>(s, src, dst, ch.nplanes); break;
      }
    }
  }
  if (SIGNAL) {
    /* every thread's stores -> CTA barrier -> fence -> count this CTA as done.
       The CTA that observes the full count owns the system-scope fence and the stores to the (peer) flags. */
    __syncthreads();
    if (threadIdx.x == 0) {
      /* release at GPU scope is enough here: the last CTA acquires the count at GPU scope and then releases at
         system scope, and causality order is transitive across the two (same pattern as the fused Jacobi kernel) */
      __threadfence();
      const unsigned prev = atomicAdd(done, 1u);
      if (prev == gridDim.x - 1) {
        *done = 0; // ready for the next launch on this stream
        __threadfence_system();
        for (unsigned k = 0; k < sig.n; ++k) st_relaxed_sys(sig.flag[k], epoch);
      }
    }
  }
}

__global__ void wait_flags_kernel(WaitSet ws, unsigned epoch, unsigned *errWord, unsigned long long timeoutNs) {
  const unsigned k = threadIdx.x;
  if (k >= ws.n) return;
  const unsigned long long t0 = globaltimer_ns();
  unsigned backoff = 32;
  while (int(ld_acquire_sys(ws.flag[k]) - epoch) < 0) {
    __nanosleep(backoff);
    if (backoff < 1024) backoff <<= 1;
    if (globaltimer_ns() - t0 > timeoutNs) {
      atomicCAS(errWord, 0u, 1u + k);
      __threadfence_system();
      return;
    }
  }
}

/* a few boxes whose descriptors arrive as kernel parameters: blockIdx.y = box, blockIdx.x strides over row groups */
__global__ void __launch_bounds__(kThreads) box_copy_batch_kernel(const __grid_constant__ SegBatch batch, int parity, unsigned rowsPerCta) {
  const CopySeg &s = batch.seg[blockIdx.y];
  /* work items = (plane, group of rowsPerCta rows inside the plane); one division per item, none per element */
  const unsigned groupsPerPlane = (s.ny + rowsPerCta - 1) / rowsPerCta;
  const unsigned items = groupsPerPlane * s.nz;
  for (unsigned it = blockIdx.x; it < items; it += gridDim.x) {
    const unsigned z = it / groupsPerPlane, y0 = (it - z * groupsPerPlane) * rowsPerCta;
    const unsigned n = min(rowsPerCta, s.ny - y0);
    const char *src = s.src[parity] + (long long)z * s.srcPitchZ + (long long)y0 * s.srcPitchY;
    char *dst = s.dst[parity] + (long long)z * s.dstPitchZ + (long long)y0 * s.dstPitchY;
    switch (s.vecLog2) {
    case 4: copy_rows<16>(s, src, dst, n); break;
    case 3: copy_rows<8>(s, src, dst, n); break;
    case 2: copy_rows<4>(s, src, dst, n); break;
    case 1: copy_rows<2>(s, src, dst, n); break;
    default: copy_rows<1>(s, src, dst, n); break;
    }
  }
}

__global__ void signal_flags_kernel(SignalSet sig, unsigned epoch) {
  if (threadIdx.x < sig.n) {
    __threadfence_system();
    st_relaxed_sys(sig.flag[threadIdx.x], epoch);
  }
}

unsigned log2_floor_pow2(unsigned long long x) {
  unsigned l = 0;
  while (l < 4 && (x % (2ull << l)) == 0) ++l;
  return l;
}

} // namespace

CopyPlan::~CopyPlan() {
  if (dSegs_ || dChunks_ || dDone_) {
    int prev = 0;
    if (cudaGetDevice(&prev) == cudaSuccess) {
      cudaSetDevice(dev_);
      cudaFree(dSegs_);
      cudaFree(dChunks_);
      cudaFree(dDone_);
      cudaSetDevice(prev);
    }
  }
}

void CopyPlan::swap(CopyPlan &o) noexcept {
  segs_.swap(o.segs_);
  chunks_.swap(o.chunks_);
  std::swap(bytes_, o.bytes_);
  std::swap(dSegs_, o.dSegs_);
  std::swap(dChunks_, o.dChunks_);
  std::swap(dDone_, o.dDone_);
  std::swap(dev_, o.dev_);
  optional_.swap(o.optional_);
  remote_.swap(o.remote_);
  std::swap(nRequiredChunks_, o.nRequiredChunks_);
}

void CopyPlan::clear() {
  CopyPlan empty;
  swap(empty);
}

CopySeg make_seg(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext) {
  const long long es = (long long)src.elemSize;
  CopySeg s;
  for (int p = 0; p < 2; ++p) {
    s.src[p] = src.base[p] + srcPos.z * src.pitchZ + srcPos.y * src.pitchY + srcPos.x * es;
    s.dst[p] = dst.base[p] + dstPos.z * dst.pitchZ + dstPos.y * dst.pitchY + dstPos.x * es;
  }
  s.srcPitchY = src.pitchY;
  s.srcPitchZ = src.pitchZ;
  s.dstPitchY = dst.pitchY;
  s.dstPitchZ = dst.pitchZ;
  s.rowBytes = unsigned(ext.x * es);
  s.ny = unsigned(ext.y);
  s.nz = unsigned(ext.z);
  /* widest power-of-two (<= 16 B) dividing every address any row can start at, and the row length */
  unsigned long long a = s.rowBytes;
  for (int p = 0; p < 2; ++p) a |= (unsigned long long)(uintptr_t)s.src[p] | (unsigned long long)(uintptr_t)s.dst[p];
  if (s.ny > 1) a |= (unsigned long long)src.pitchY | (unsigned long long)dst.pitchY;
  if (s.nz > 1) a |= (unsigned long long)src.pitchZ | (unsigned long long)dst.pitchZ;
  s.vecLog2 = log2_floor_pow2(a);
  /* sector mode: thin rows whose destination touches row padding on the side(s) where it is not sector aligned */
  s.padLoWords = s.totalWords = 0;
  const bool pitchesOk = ((dst.pitchY | dst.pitchZ) & 31) == 0 && ((src.pitchY | src.pitchZ) & 3) == 0;
  if (pitchesOk && (es % 4) == 0 && s.rowBytes <= 56) {
    bool ok = true;
    unsigned lo = 0, hi = 0;
    for (int p = 0; p < 2 && ok; ++p) {
      const unsigned long long start = (unsigned long long)(uintptr_t)s.dst[p], end = start + s.rowBytes;
      const unsigned l = unsigned(start & 31), h = unsigned((32 - (end & 31)) & 31);
      if (((uintptr_t)s.src[p] & 3) != 0) ok = false;
      if (p == 0) {
        lo = l;
        hi = h;
      } else if (l != lo || h != hi) {
        ok = false;
      }
    }
    if (lo && !(dstPos.x == 0 && (long long)lo <= dst.padLoBytes)) ok = false;
    if (hi && !(dst.rowElems > 0 && dstPos.x + ext.x == dst.rowElems && (long long)hi <= dst.padHiBytes)) ok = false;
    if (ok && (lo || hi) && lo + s.rowBytes + hi <= 64) {
      s.padLoWords = (unsigned short)(lo / 4);
      s.totalWords = (unsigned short)((lo + s.rowBytes + hi) / 4);
    }
  }
  /* thin rows (<= 16 bytes) out of a padded array whose aligned windows may be read: same word offset in every row */
  s.srcWin = 0;
  if (src.windowSafe && (es % 4) == 0 && s.rowBytes <= 16 && ((src.pitchY | src.pitchZ) & 15) == 0 && (((uintptr_t)s.src[0] | (uintptr_t)s.src[1]) & 3) == 0 &&
      (((uintptr_t)s.src[0] ^ (uintptr_t)s.src[1]) & 15) == 0)
    s.srcWin = 1u + unsigned(((uintptr_t)s.src[0] >> 2) & 3u);
  /* thin rows into a dense destination: gather through shared memory — always when the rows can be fetched by window
     (the stores then leave as coalesced 16-byte runs instead of one 4-byte store per word), else only across NVLink */
  s.gatherDense = 0;
  if ((dst.acrossLink || s.srcWin) && s.totalWords == 0 && (es % 4) == 0 && s.rowBytes <= 64 && dst.pitchY == (long long)s.rowBytes &&
      (s.nz == 1 || dst.pitchZ == (long long)s.rowBytes * s.ny) && ((src.pitchY | src.pitchZ) & 3) == 0) {
    bool ok = true;
    for (int p = 0; p < 2; ++p) ok = ok && (((uintptr_t)s.src[p] | (uintptr_t)s.dst[p]) & 3) == 0;
    s.gatherDense = ok ? 1u : 0u;
  }
  return s;
}

void CopyPlan::add(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext) {
  if (ext.x <= 0 || ext.y <= 0 || ext.z <= 0) return;
  const CopySeg s = make_seg(src, srcPos, dst, dstPos, ext);
  segs_.push_back(s);
  optional_.push_back(0);
  remote_.push_back(dst.acrossLink ? 1 : 0);
  bytes_ += size_t(s.rowBytes) * s.ny * s.nz;
}

void CopyPlan::add_optional(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext) {
  const size_t before = segs_.size();
  add(src, srcPos, dst, dstPos, ext);
  if (segs_.size() > before) optional_.back() = 1;
}

void CopyPlan::finalize(int dev, size_t chunkBytes) {
  dev_ = dev;
  build_chunks(chunkBytes);
  upload();
}

void CopyPlan::build_chunks(size_t chunkBytes) {
  chunks_.clear();
  nRequiredChunks_ = 0;
  /* chunks of the optional segments go last, so that leaving them out is a shorter chunk table */
  for (int pass = 0; pass < 2; ++pass) {
  if (pass == 1) nRequiredChunks_ = unsigned(chunks_.size());
  /* Chunks whose destination is another GPU are spread evenly among the local ones: a launch walks the table front to
     back, and the two kinds are bound by different things — NVLink for the first, the DRAM's page-open rate for thin
     local rows — so each should be in flight during the whole launch instead of one after the other. */
  std::vector<CopyChunk> link, local;
  for (size_t si = 0; si < segs_.size(); ++si) {
    if (int(optional_[si]) != pass) continue;
    const CopySeg &s = segs_[si];
    std::vector<CopyChunk> &out = remote_[si] ? link : local;
    /* thin rows cost a sector each way whatever their length: count at least 32 bytes per row when sizing chunks */
    const unsigned per = unsigned(std::min<size_t>(32768, std::max<size_t>(1, chunkBytes / std::max(32u, s.rowBytes))));
    if (s.ny >= per) {
      for (unsigned z = 0; z < s.nz; ++z)
        for (unsigned y = 0; y < s.ny; y += per)
          out.push_back(CopyChunk{unsigned(si), z, y, (unsigned short)std::min(per, s.ny - y), (unsigned short)1});
    } else { // few rows per plane (edges, corners): a chunk takes several whole planes
      const unsigned planesPer = std::max(1u, per / s.ny);
      for (unsigned z = 0; z < s.nz; z += planesPer)
        out.push_back(CopyChunk{unsigned(si), z, 0, (unsigned short)s.ny, (unsigned short)std::min(planesPer, s.nz - z)});
    }
  }
  {
    size_t a = 0, b = 0;
    const size_t na = link.size(), nb = local.size();
    while (a < na || b < nb) {
      /* the kind that is further behind its share goes next */
      if (b >= nb || (a < na && a * nb <= b * na))
        chunks_.push_back(link[a++]);
      else
        chunks_.push_back(local[b++]);
    }
  }
  }
}

void CopyPlan::upload() {
  if (segs_.empty()) return;
  int prev = 0;
  CUDA_RUNTIME(cudaGetDevice(&prev));
  CUDA_RUNTIME(cudaSetDevice(dev_));
  if (dSegs_) CUDA_RUNTIME(cudaFree(dSegs_));
  if (dChunks_) CUDA_RUNTIME(cudaFree(dChunks_));
  if (!dDone_) {
    CUDA_RUNTIME(cudaMalloc(&dDone_, sizeof(unsigned)));
    CUDA_RUNTIME(cudaMemset(dDone_, 0, sizeof(unsigned)));
  }
  CUDA_RUNTIME(cudaMalloc(&dSegs_, segs_.size() * sizeof(CopySeg)));
  CUDA_RUNTIME(cudaMalloc(&dChunks_, chunks_.size() * sizeof(CopyChunk)));
  CUDA_RUNTIME(cudaMemcpy(dSegs_, segs_.data(), segs_.size() * sizeof(CopySeg), cudaMemcpyHostToDevice));
  CUDA_RUNTIME(cudaMemcpy(dChunks_, chunks_.data(), chunks_.size() * sizeof(CopyChunk), cudaMemcpyHostToDevice));
  CUDA_RUNTIME(cudaSetDevice(prev));
}

unsigned CopyPlan::grid_size(bool withOptional) const {
  /* 148 SMs x 8 resident 256-thread CTAs is a full wave on B200; small plans use one CTA per chunk */
  return unsigned(std::min<size_t>(withOptional ? chunks_.size() : nRequiredChunks_, 148 * 8));
}

void CopyPlan::launch(int parity, cudaStream_t stream) const {
  if (chunks_.empty()) return;
  SignalSet none;
  none.n = 0;
  box_copy_kernel<false><<<grid_size(), kThreads, 0, stream>>>(dSegs_, dChunks_, unsigned(chunks_.size()), parity, dDone_, none, 0u);
  CUDA_RUNTIME(cudaGetLastError());
}

void CopyPlan::launch_and_signal(int parity, const SignalSet &sig, unsigned epoch, cudaStream_t stream, bool withOptional) const {
  const unsigned n = withOptional ? unsigned(chunks_.size()) : nRequiredChunks_;
  if (n == 0) {
    /* nothing to move, but neighbours still wait for the flag */
    launch_signal_flags(sig, epoch, stream);
    return;
  }
  box_copy_kernel<true><<<grid_size(withOptional), kThreads, 0, stream>>>(dSegs_, dChunks_, n, parity, dDone_, sig, epoch);
  CUDA_RUNTIME(cudaGetLastError());
}

void launch_box_copy(const SegBatch &batch, int parity, cudaStream_t stream) {
  if (batch.n == 0) return;
  unsigned maxItems = 1, minRowBytes = 0xffffffffu;
  for (unsigned i = 0; i < batch.n; ++i) minRowBytes = std::min(minRowBytes, std::max(32u, batch.seg[i].rowBytes));
  const unsigned rowsPerCta = std::max(1u, (32u * 1024u) / minRowBytes);
  for (unsigned i = 0; i < batch.n; ++i) maxItems = std::max(maxItems, (batch.seg[i].ny + rowsPerCta - 1) / rowsPerCta * batch.seg[i].nz);
  const unsigned gx = std::min(maxItems, 148u * 8u);
  box_copy_batch_kernel<<<dim3(gx, batch.n, 1), kThreads, 0, stream>>>(batch, parity, rowsPerCta);
  CUDA_RUNTIME(cudaGetLastError());
}

void launch_signal_flags(const SignalSet &sig, unsigned epoch, cudaStream_t stream) {
  if (sig.n == 0) return;
  signal_flags_kernel<<<1, 32, 0, stream>>>(sig, epoch);
  CUDA_RUNTIME(cudaGetLastError());
}

void launch_wait_flags(const WaitSet &ws, unsigned epoch, unsigned *errWord, unsigned long long timeoutNs, cudaStream_t stream) {
  if (ws.n == 0) return;
  wait_flags_kernel<<<1, 32, 0, stream>>>(ws, epoch, errWord, timeoutNs);
  CUDA_RUNTIME(cudaGetLastError());
}

} // namespace stencil
