/* AstarothSim engine and kernels (N4/N5 radius-3 multi-field stencil, N9 init).
   Replaces reference kernels K10/K11 (bin/astaroth_sim.cu:14-83). See include/stencil/astaroth.hpp.

   Kernel shape: a CTA is 32 lanes x 8 rows, a lane owns one 16-byte vector of cells (4 fp32 / 2 fp64) and marches
   along z with the seven centre values z-3..z+3 rotating through registers, so the z direction costs one load per
   plane; x and y neighbours are 16-byte loads that hit L1 (rows are 128-byte aligned). All fields go in one launch
   (blockIdx.z = field x z-chunk). That is the plain kernel (thin regions, Average7); the Star19 hot path is the
   TMA-pipelined kernel further down. Schedules: the reference's interior / exchange() / exterior with a host
   synchronisation per iteration, or (default with the TMA kernel) exchange_async() + one whole-region launch per
   subdomain chained by events, with no host synchronisation in the loop. Folding the push into the kernel as the
   Jacobi engine does is the next step. */
#include "stencil/nvtx.hpp"
#include "stencil/astaroth.hpp"

#include <cmath>
#include <cstring>
#include <limits>
#include <set>

#include "stencil/copy_plan.hpp"
#include "stencil/detail/tma.cuh"

namespace stencil {

namespace {

constexpr int kMaxFields = 16, kLanes = 32, kRows = 8, kR = 3;

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int d) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != d) CUDA_RUNTIME(cudaSetDevice(d));
  }
  ~DevGuard() { cudaSetDevice(prev); }
};

template <typename T> struct Vec;
template <> struct Vec<float> {
  typedef float4 type;
  static constexpr int n = 4;
};
template <> struct Vec<double> {
  typedef double2 type;
  static constexpr int n = 2;
};

template <typename T> struct FieldParams {
  const T *src[kMaxFields]; // compute cell (0,0,0) of curr, per field
  T *dst[kMaxFields];
  long long pitchX, pitchXY;
  int lox, loy, loz, hix, hiy, hiz; // compute coordinates, half open
  int tileZ, nChunks;
  T dt;
};

/* 6th-order second-derivative weights: c0 = -49/18 per axis, then 3/2, -3/20, 1/90 */
template <typename T> __host__ __device__ inline T star19(const T c, const T xm[3], const T xp[3], const T ym[3], const T yp[3], const T zm[3], const T zp[3], T dt) {
  const T c1 = T(1.5), c2 = T(-0.15), c3 = T(1.0 / 90.0), c0 = T(-49.0 / 18.0);
  T lap = T(3) * c0 * c;
  lap += c1 * (((((xm[0] + xp[0]) + ym[0]) + yp[0]) + zm[0]) + zp[0]);
  lap += c2 * (((((xm[1] + xp[1]) + ym[1]) + yp[1]) + zm[1]) + zp[1]);
  lap += c3 * (((((xm[2] + xp[2]) + ym[2]) + yp[2]) + zm[2]) + zp[2]);
  return c + dt * lap;
}

/* the reference's stand-in: (-x + -y + -z + +x + +y + +z) / 6 in that order */
template <typename T> __host__ __device__ inline T average7(T xm, T ym, T zm, T xp, T yp, T zp) {
  T v = 0;
  v += xm;
  v += ym;
  v += zm;
  v += xp;
  v += yp;
  v += zp;
  return v / T(6);
}

template <typename T, bool STAR> __global__ void __launch_bounds__(kLanes *kRows) astaroth_kernel(const __grid_constant__ FieldParams<T> P) {
  typedef typename Vec<T>::type V;
  constexpr int N = Vec<T>::n;
  const int field = blockIdx.z / P.nChunks, chunk = blockIdx.z % P.nChunks;
  const int x0 = (P.lox / N) * N + (blockIdx.x * kLanes + threadIdx.x) * N;
  const int y = P.loy + blockIdx.y * kRows + threadIdx.y;
  const int z0 = P.loz + chunk * P.tileZ, z1 = min(z0 + P.tileZ, P.hiz);
  if (y >= P.hiy || x0 >= P.hix || x0 + N <= P.lox || z0 >= z1) return;
  const T *s = P.src[field] + (long long)y * P.pitchX + x0;
  T *d = P.dst[field] + (long long)y * P.pitchX + x0;
  auto ldv = [](const T *p, T out[N]) {
    const V v = *reinterpret_cast<const V *>(p);
    std::memcpy(out, &v, sizeof(V));
  };
  /* centre column z-3 .. z+3 (STAR) or z-1 .. z+1 (average) */
  T q[7][N];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    if (STAR || (k >= 2 && k <= 4)) ldv(s + (long long)(z0 + k - 3) * P.pitchXY, q[k]);
  }
  for (int z = z0; z < z1; ++z) {
    const T *sp = s + (long long)z * P.pitchXY;
    T out[N];
    if (STAR) {
      /* x neighbours: the vectors left and right of mine (and one further for fp64) */
      T xl[2 * N], xr[2 * N];
      ldv(sp - N, xl + N);
      ldv(sp + N, xr);
      if (N < 3) {
        ldv(sp - 2 * N, xl);
        ldv(sp + 2 * N, xr + N);
      }
      T ym[3][N], yp[3][N];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        ldv(sp - (long long)(k + 1) * P.pitchX, ym[k]);
        ldv(sp + (long long)(k + 1) * P.pitchX, yp[k]);
      }
#pragma unroll
      for (int i = 0; i < N; ++i) {
        /* cells in x order around cell i: window = [xl | q[3] | xr] */
        T win[4 * N + N];
#pragma unroll
        for (int j = 0; j < 2 * N; ++j) win[j] = xl[j];
#pragma unroll
        for (int j = 0; j < N; ++j) win[2 * N + j] = q[3][j];
#pragma unroll
        for (int j = 0; j < 2 * N; ++j) win[3 * N + j] = xr[j];
        const int c = 2 * N + i;
        const T xm[3] = {win[c - 1], win[c - 2], win[c - 3]}, xp[3] = {win[c + 1], win[c + 2], win[c + 3]};
        const T ymv[3] = {ym[0][i], ym[1][i], ym[2][i]}, ypv[3] = {yp[0][i], yp[1][i], yp[2][i]};
        const T zm[3] = {q[2][i], q[1][i], q[0][i]}, zp[3] = {q[4][i], q[5][i], q[6][i]};
        out[i] = star19<T>(q[3][i], xm, xp, ymv, ypv, zm, zp, P.dt);
      }
    } else {
      T l[N], r[N], ym[N], yp[N];
      ldv(sp - N, l);
      ldv(sp + N, r);
      ldv(sp - P.pitchX, ym);
      ldv(sp + P.pitchX, yp);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const T xm = i == 0 ? l[N - 1] : q[3][i - 1], xp = i == N - 1 ? r[0] : q[3][i + 1];
        out[i] = average7<T>(xm, ym[i], q[2][i], xp, yp[i], q[4][i]);
      }
    }
    T *dp = d + (long long)z * P.pitchXY;
    if (x0 >= P.lox && x0 + N <= P.hix) {
      V v;
      std::memcpy(&v, out, sizeof(V));
      *reinterpret_cast<V *>(dp) = v;
    } else {
#pragma unroll
      for (int i = 0; i < N; ++i)
        if (x0 + i >= P.lox && x0 + i < P.hix) dp[i] = out[i];
    }
    /* rotate the centre column and fetch the next plane */
#pragma unroll
    for (int k = (STAR ? 0 : 2); k < (STAR ? 6 : 4); ++k)
#pragma unroll
      for (int i = 0; i < N; ++i) q[k][i] = q[k + 1][i];
    if (z + 1 < z1) ldv(s + (long long)(z + 1 + (STAR ? 3 : 1)) * P.pitchXY, q[STAR ? 6 : 4]);
  }
}

/* ---------------- TMA-pipelined radius-3 star kernel ----------------
   Same structure as the Jacobi kernel (src/jacobi.cu): a tile is 32 lanes x 16 rows, a lane owns one 16-byte vector
   of cells (N = 4 fp32 / 2 fp64), thread 0 streams (32N + 8) x (16 + 6) input planes of ONE field into a 6-stage
   shared-memory ring with cp.async.bulk.tensor.3d, eight warps (two rows each) consume. The z direction never goes
   back to memory: each arriving plane adds its "above" contribution to the three pending outputs and its centre
   stays in registers for the three "below" contributions that follow, so a plane is read from shared memory once. */
constexpr int kTRows = 16, kTMargin = 4, kTStages = 6, kTLag = 2, kTThreads = 256;

template <typename T> struct TmaGeom {
  static constexpr int N = Vec<T>::n;
  static constexpr int tileX = 32 * N;
  static constexpr int boxX = tileX + 2 * kTMargin;
  static constexpr int boxY = kTRows + 2 * kR;
  static constexpr int planeBytes = boxX * boxY * int(sizeof(T));
  static constexpr int stageBytes = (planeBytes + 127) / 128 * 128;
  static constexpr int smemBytes = kTStages * stageBytes + 2 * kTStages * 8;
};

/* Fused y / z face pushes: tiles on a y or z face of the subdomain also store their rows into
   the neighbour's `next` halo — peer pointers over NVLink, or plain stores when the neighbour shares the GPU — and
   the kernels of neighbouring GPUs hand-shake through epoch flags exactly as the fused Jacobi kernel does
   (src/jacobi.cu): boundary CTAs wait for the neighbours' flags of the previous iteration before reading halos,
   count themselves done after a GPU-scope fence, and the last one raises the neighbours' flags behind one
   system-scope fence. A 19-point star reads face halos only, so in steady state no exchange() is left. */
enum PushFace { kYM = 0, kYP = 1, kZM = 2, kZP = 3 };
constexpr int kFusedFlagBase = 64; // word offset of the fused-iteration flags inside a control block (as in jacobi.cu)

template <typename T> struct PushParams {
  T *nb[4][kMaxFields];      // my compute cell (x,y,z) on face f lands at nb[f][field] + z*nbPitchXY[f] + y*nbPitchX[f] + x
  long long nbPitchX[4], nbPitchXY[4];
  unsigned *nbFlag[4];       // neighbour's flag to raise (nullptr: ordered by streams on one GPU)
  const unsigned *myFlag[4]; // my flag the neighbour raises
  unsigned faceMask, flagMask;
  unsigned nBoundary; // CTAs that take part in the hand-shake
  unsigned *done, *err;
  unsigned long long timeoutNs;
  int sx, sy, sz;
};

__device__ __forceinline__ void st_flag_sys(unsigned *p, unsigned v) {
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

/* Boundary tiles only, after their plane loop: copy the rows of this tile that lie on y / z faces from my `next`
   buffer (written a moment ago by this CTA) to the matching halo rows of the neighbours' `next` buffers.
   One warp per row, 16-byte vectors when both sides are aligned. Loads bypass L1 (ld.global.cg). */
template <typename T>
__device__ __noinline__ void push_tile_faces(const PushParams<T> &Q, int field, unsigned faces, int xlo, int ylo, int z0, int z1, const T *mine,
                                             long long pitchX, long long pitchXY) {
  typedef typename Vec<T>::type V;
  constexpr int N = Vec<T>::n;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = kTThreads / 32;
  const int xhi = min(xlo + 32 * N, Q.sx), yhi = min(ylo + kTRows, Q.sy);
  for (int f = 0; f < 4; ++f) {
    if (!((faces >> f) & 1u)) continue;
    int ya = ylo, yb = yhi, za = z0, zb = z1;
    if (f == kYM) yb = min(yb, kR);
    if (f == kYP) ya = max(ya, Q.sy - kR);
    if (f == kZM) zb = min(zb, kR);
    if (f == kZP) za = max(za, Q.sz - kR);
    const int ny = yb - ya, rows = ny * (zb - za);
    T *nb = Q.nb[f][field];
    const long long npx = Q.nbPitchX[f], npxy = Q.nbPitchXY[f];
    for (int r = warp; r < rows; r += nwarps) {
      const int z = za + r / ny, y = ya + r % ny;
      const T *src = mine + (long long)z * pitchXY + (long long)y * pitchX;
      T *dst = nb + (long long)z * npxy + (long long)y * npx;
      const bool vec = ((reinterpret_cast<unsigned long long>(dst + xlo) | reinterpret_cast<unsigned long long>(src + xlo)) & (sizeof(V) - 1)) == 0;
      if (vec) {
        const int x = xlo + lane * N;
        if (x + N <= xhi) {
          *reinterpret_cast<V *>(dst + x) = __ldcg(reinterpret_cast<const V *>(src + x));
        } else {
          for (int i = x; i < xhi; ++i) dst[i] = __ldcg(src + i);
        }
      } else {
        for (int x = xlo + lane; x < xhi; x += 32) dst[x] = __ldcg(src + x);
      }
    }
  }
}

template <typename T> __device__ __noinline__ void wait_neighbours(const PushParams<T> &Q, unsigned faces, unsigned epoch) {
  if (threadIdx.x < 4 && ((faces >> threadIdx.x) & 1u)) {
    const unsigned *f = Q.myFlag[threadIdx.x];
    if (f) {
      const unsigned long long t0 = globaltimer_ns();
      while (int(ld_acquire_sys(f) - (epoch - 1)) < 0) {
        __nanosleep(64);
        if (globaltimer_ns() - t0 > Q.timeoutNs) {
          atomicCAS(Q.err, 0u, 200u + threadIdx.x);
          break;
        }
      }
    }
  }
}

template <typename T> __device__ __noinline__ void signal_neighbours(const PushParams<T> &Q, unsigned epoch) {
  __threadfence();
  const unsigned prev = atomicAdd(Q.done, 1u);
  if (prev == Q.nBoundary - 1) {
    *Q.done = 0;
    __threadfence_system();
    for (int f = 0; f < 4; ++f)
      if (Q.nbFlag[f]) st_flag_sys(Q.nbFlag[f], epoch);
  }
}

/* which y / z faces does tile (ty, z-chunk [z0,z1)) own cells on? (host and device agree through this one function) */
__host__ __device__ inline unsigned tile_faces(int ty, int z0, int z1, int sy, int sz, unsigned faceMask) {
  unsigned f = 0;
  const int ylo = ty * 16, yhi = ylo + 16;
  if (ylo < kR) f |= 1u << kYM;
  if (yhi > sy - kR && ylo < sy) f |= 1u << kYP;
  if (z0 < kR) f |= 1u << kZM;
  if (z1 > sz - kR) f |= 1u << kZP;
  return f & faceMask;
}

template <typename T> struct TmaParams {
  T *dst[kMaxFields];      // compute cell (0,0,0) of next, per field
  const CUtensorMap *maps; // device array: one map per field over the CURRENT curr buffers
  long long pitchX, pitchXY;
  int lox, loy, loz, hix, hiy, hiz; // cells to update, compute coordinates, half open
  int tileZ, tilesX, tilesY, nChunks;
  int tx0, ty0;            // first tile column / row of the region (tile grid is anchored at compute cell 0)
  int x0pad;               // index of compute x = 0 inside a padded row (TMA coordinates start at the row start)
  const PushParams<T> *push; // device pointer, nullptr: no y / z pushes
  unsigned pushFaceMask, pushFlagMask; // copies of push->faceMask / flagMask (decided before anything is loaded)
  unsigned epoch;                      // iteration being computed (1-based)
  int sy, sz;                          // compute extent of the subdomain (for the face tests)
  int wrapX;               // > 0: the subdomain is its own +-x neighbour and is wrapX cells wide; the kernel then also
                           // writes its first / last three cells of every row into the opposite x halo of `next`
  T dt;
};

template <typename T> __global__ void __launch_bounds__(kTThreads, 2) astaroth_tma_kernel(const __grid_constant__ TmaParams<T> P) {
  typedef TmaGeom<T> G;
  typedef typename Vec<T>::type V;
  constexpr int N = G::N;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned long long *full = reinterpret_cast<unsigned long long *>(smem + kTStages * G::stageBytes);
  unsigned long long *empty = full + kTStages;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, field = blockIdx.y;
  int tile = blockIdx.x;
  const int tx = P.tx0 + tile % P.tilesX;
  tile /= P.tilesX;
  const int ty = P.ty0 + tile % P.tilesY, chunk = tile / P.tilesY;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTStages; ++s) {
      tma::mbar_init(&full[s], 1);
      tma::mbar_init(&empty[s], kTThreads / 32);
    }
    tma::mbar_fence_init();
  }
  const int z0 = P.loz + chunk * P.tileZ, z1 = min(z0 + P.tileZ, P.hiz);
  /* y / z faces this tile pushes to neighbours; tiles that exchange with another GPU first wait for that GPU's flags of
     the previous iteration (they are the only tiles that read halo cells it wrote) */
  const unsigned faces = P.push ? tile_faces(ty, z0, z1, P.sy, P.sz, P.pushFaceMask) : 0u;
  if (faces & P.pushFlagMask) wait_neighbours<T>(*P.push, faces & P.pushFlagMask, P.epoch);
  __syncthreads();

  const int nplanes = z1 - z0 + 2 * kR; // input planes z0-3 .. z1+2
  const CUtensorMap *map = P.maps + field;
  const int cx = P.x0pad + tx * G::tileX - kTMargin; // raw x of smem column 0
  const int cy = ty * kTRows;                        // raw row of compute row ty*16 - 3 (raw = compute + 3)
  auto issue = [&](int j) {                          // input plane j = compute plane z0 - 3 + j = raw plane z0 + j
    const int sj = j % kTStages;
    if (j >= kTStages) tma::mbar_wait(&empty[sj], ((j / kTStages) - 1) & 1);
    tma::mbar_arrive_expect_tx(&full[sj], G::planeBytes);
    tma::load_3d(smem + sj * G::stageBytes, map, cx, cy, z0 + j, &full[sj]);
  };
  if (threadIdx.x == 0)
    for (int j = 0; j < kTStages - kTLag && j < nplanes; ++j) issue(j);

  const int r0 = 2 * warp;
  const int x0 = tx * G::tileX + lane * N, y0 = ty * kTRows + r0;
  /* where the two rows of this lane go, decided once */
  T *dp[2];
  int mode[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int y = y0 + r;
    const bool in = y >= P.loy && y < P.hiy && x0 < P.hix && x0 + N > P.lox;
    mode[r] = !in ? 0 : ((x0 >= P.lox && x0 + N <= P.hix) ? 1 : 2);
    dp[r] = P.dst[field] + (long long)z0 * P.pitchXY + (long long)y * P.pitchX + x0;
  }
  /* Fused x faces (periodic wrap onto this subdomain): the lanes that own cells 0..2 or sx-3..sx-1 store their output
     vector a second time, sx cells further / earlier in the same row — that is where the opposite halo lives. The extra
     cells of the vector land in row padding. fp32 vectors are half a 32-byte sector: the other half (padding only) is
     written as well so that no sector is left partially written. */
  int qdelta = 0, qwide = 0;
  const bool qvec = P.wrapX % N == 0;
  if (P.wrapX > 0) {
    if (x0 < kR) {
      qdelta = P.wrapX;
      qwide = (N == 4 && x0 == 0 && qvec) ? N : 0;
    } else if (x0 + N > P.wrapX - kR && x0 < P.wrapX) {
      qdelta = -P.wrapX;
      qwide = (N == 4 && x0 + N == P.wrapX && qvec) ? -N : 0;
    }
  }
  const T c1 = T(1.5), c2 = T(-0.15), c3 = T(1.0 / 90.0), c0x3 = T(3) * T(-49.0 / 18.0);
  /* pending sums for the planes 1, 2, 3 behind the newest one, and the centres of those planes */
  T a1[2][N], a2[2][N], a3[2][N], h1[2][N], h2[2][N], h3[2][N];
#pragma unroll
  for (int r = 0; r < 2; ++r)
#pragma unroll
    for (int i = 0; i < N; ++i) a1[r][i] = a2[r][i] = a3[r][i] = h1[r][i] = h2[r][i] = h3[r][i] = T(0);

  int s = 0;
  unsigned phase = 0;
  for (int j = 0; j < nplanes; ++j) {
    tma::mbar_wait(&full[s], phase);
    const T *pl = reinterpret_cast<const T *>(smem + s * G::stageBytes);
    /* smem row q holds compute row ty*16 + q - 3; smem column kTMargin + k holds compute x = tx*tileX + k */
    const T *base = pl + (r0 + kR) * G::boxX + kTMargin + lane * N;
    T col[8][N]; // rows r0-3 .. r0+4 at my x: col[3], col[4] are my two centre rows
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const V v = *reinterpret_cast<const V *>(base + (q - 3) * G::boxX);
      memcpy(col[q], &v, sizeof(V));
    }
    const bool inPlane = j >= kR && j < nplanes - kR; // this plane is an output plane: it needs its in-plane sum
    T an[2][N];
    if (inPlane) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        /* x window of row r: cells x0 - kTMargin .. x0 + N + kTMargin - 1 */
        T win[N + 2 * kTMargin];
        const T *row = base + r * G::boxX;
#pragma unroll
        for (int v = 0; v < (N + 2 * kTMargin) / N; ++v) {
          if (v * N == kTMargin) {
#pragma unroll
            for (int i = 0; i < N; ++i) win[kTMargin + i] = col[3 + r][i];
          } else {
            const V t = *reinterpret_cast<const V *>(row + v * N - kTMargin);
            memcpy(win + v * N, &t, sizeof(V));
          }
        }
#pragma unroll
        for (int i = 0; i < N; ++i) {
          const int c = kTMargin + i;
          T lap = c0x3 * win[c];
          lap += c1 * (((win[c - 1] + win[c + 1]) + col[3 + r - 1][i]) + col[3 + r + 1][i]);
          lap += c2 * (((win[c - 2] + win[c + 2]) + col[3 + r - 2][i]) + col[3 + r + 2][i]);
          lap += c3 * (((win[c - 3] + win[c + 3]) + col[3 + r - 3][i]) + col[3 + r + 3][i]);
          /* contributions of the three planes below, whose centres are still in registers */
          lap += c1 * h1[r][i];
          lap += c2 * h2[r][i];
          lap += c3 * h3[r][i];
          an[r][i] = lap;
        }
      }
    } else {
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int i = 0; i < N; ++i) an[r][i] = T(0);
    }
    /* the slot can be refilled: everything needed from it is in registers */
    __syncwarp();
    if (lane == 0) tma::mbar_arrive(&empty[s]);
    if (threadIdx.x == 0 && j + kTStages - kTLag < nplanes) issue(j + kTStages - kTLag);
    if (++s == kTStages) {
      s = 0;
      phase ^= 1u;
    }
    /* this plane is "above" the three pending ones; the oldest of them is now complete */
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      T out[N];
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const T c = col[3 + r][i];
        out[i] = h3[r][i] + P.dt * (a3[r][i] + c3 * c);
        a3[r][i] = a2[r][i] + c2 * c;
        a2[r][i] = a1[r][i] + c1 * c;
        a1[r][i] = an[r][i];
        h3[r][i] = h2[r][i];
        h2[r][i] = h1[r][i];
        h1[r][i] = c;
      }
      if (j >= 2 * kR) { // output plane z0 + j - 6
        if (mode[r] == 1) {
          V v;
          memcpy(&v, out, sizeof(V));
          *reinterpret_cast<V *>(dp[r]) = v;
          if (qdelta) {
            if (qvec) {
              *reinterpret_cast<V *>(dp[r] + qdelta) = v;
              if (qwide) *reinterpret_cast<V *>(dp[r] + qdelta + qwide) = v; // padding cells: any value will do
            } else { // row length not a multiple of the vector: element stores, face cells only
#pragma unroll
              for (int i = 0; i < N; ++i)
                if (x0 + i < kR || x0 + i >= P.wrapX - kR) dp[r][qdelta + i] = out[i];
            }
          }
        } else if (mode[r] == 2) {
#pragma unroll
          for (int i = 0; i < N; ++i)
            if (x0 + i >= P.lox && x0 + i < P.hix) {
              dp[r][i] = out[i];
              if (qdelta && (x0 + i < kR || x0 + i >= P.wrapX - kR)) dp[r][qdelta + i] = out[i];
            }
        }
        dp[r] += P.pitchXY;
      }
    }
  }
  if (faces) {
    /* The plane loop above is untouched by the fusion. A boundary tile copies the face rows it has just written
       (still in L2) into the neighbours' halos here, whole CTA, one warp per row. */
    __syncthreads();
    push_tile_faces<T>(*P.push, field, faces, tx * G::tileX, ty * kTRows, z0, z1, P.dst[field], P.pitchX, P.pitchXY);
    if (faces & P.pushFlagMask) {
      __syncthreads();
      if (threadIdx.x == 0) signal_neighbours<T>(*P.push, P.epoch);
    }
  }
}

template <typename T>
__global__ void astaroth_init_kernel(T *logical0, long long pitchX, long long pitchXY, Dim3 raw, Dim3 haloOrigin, double period) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= raw.x || y >= raw.y) return;
  for (int z = blockIdx.z; z < raw.z; z += gridDim.z) {
    const bool inside = x >= kR && y >= kR && z >= kR && x < raw.x - kR && y < raw.y - kR && z < raw.z - kR;
    T v = T(-10);
    if (inside) {
      const double w = 2 * 3.14159 / period;
      v = T(sin(w * double(haloOrigin.x + x) + w * double(haloOrigin.y + y) + w * double(haloOrigin.z + z)));
    }
    logical0[(long long)z * pitchXY + (long long)y * pitchX + x] = v;
  }
}

} // namespace

struct AstarothSim::Impl {
  std::vector<RcStream> streams;
  std::vector<CUtensorMap *> maps[2]; // per subdomain: device array of one tensor map per field, per curr parity
  std::vector<Rect3> interior;
  std::vector<std::vector<Rect3>> exterior;
  double exchangeSeconds = 0;
  /* free-running schedule: no host synchronisation inside the loop (exchange_async + events) */
  bool freeRun = false;
  /* the kernel writes the +-x halos of `next` itself (subdomains that wrap onto themselves in x) and exchange() leaves
     those faces out */
  bool fusedX = false;
  bool xHalosFromKernel = false; // the previous step's kernel filled the x halos of what is now `curr`
  /* y / z faces pushed by the kernel too (with fusedX: no exchange() left in steady state) */
  bool fusedYZ = false;
  struct Push {
    void *dev[2] = {nullptr, nullptr}; // PushParams<T> per curr parity, in device memory
    unsigned faceMask = 0, flagMask = 0;
    unsigned *done = nullptr;
    std::vector<int> sameDevDeps; // local subdomains on the same GPU that push into me / I push into
    WaitSet finalWait;
  };
  std::vector<Push> push;
  unsigned iter = 0;       // iterations launched (the epoch published in the flags)
  unsigned *err = nullptr; // host-mapped error word for flag time-outs
  std::vector<cudaEvent_t> computeDone;
  std::vector<bool> computeRecorded;
};

namespace {
const Dim3 kPushDir[4] = {Dim3(0, -1, 0), Dim3(0, 1, 0), Dim3(0, 0, -1), Dim3(0, 0, 1)};

/* one PushParams<T> (device memory) for local subdomain di with `curr` = buffer `parity` */
template <typename T>
void *build_push(DistributedDomain &dd, size_t di, int parity, int fields, const NeighborInfo (*nbs)[kMaxFields], unsigned faceMask, unsigned flagMask,
                 unsigned nBoundary, unsigned *done, unsigned *err) {
  LocalDomain &ld = dd.domains()[di];
  PushParams<T> Q;
  std::memset(&Q, 0, sizeof(Q));
  const Dim3 sz = ld.size();
  for (int f = 0; f < 4; ++f) {
    if (!((faceMask >> f) & 1u)) continue;
    for (int q = 0; q < fields; ++q) {
      const NeighborInfo &n = nbs[f][q];
      const long long npx = n.view.pitchY / (long long)sizeof(T), npxy = n.view.pitchZ / (long long)sizeof(T);
      T *interior = reinterpret_cast<T *>(n.view.base[parity ^ 1]) + (long long)kR * (npxy + npx + 1); // its NEXT buffer
      long long shift = 0;
      switch (f) {
      case kYM: shift = n.size.y * npx; break;  // my rows 0..2 -> its +y halo
      case kYP: shift = -sz.y * npx; break;     // my last rows -> its -y halo
      case kZM: shift = n.size.z * npxy; break;
      case kZP: shift = -sz.z * npxy; break;
      }
      Q.nb[f][q] = interior + shift;
      Q.nbPitchX[f] = npx;
      Q.nbPitchXY[f] = npxy;
    }
    if ((flagMask >> f) & 1u) {
      /* I raise the slot "from direction -dir" in the neighbour; the neighbour raises slot "from dir" in me */
      Q.nbFlag[f] = nbs[f][0].control + kFusedFlagBase + DirectionMap<int>::slot(kPushDir[f] * -1);
      Q.myFlag[f] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(kPushDir[f]);
    }
  }
  Q.faceMask = faceMask;
  Q.flagMask = flagMask;
  Q.nBoundary = nBoundary;
  Q.done = done;
  Q.err = err;
  Q.timeoutNs = 5000000000ull;
  Q.sx = int(sz.x);
  Q.sy = int(sz.y);
  Q.sz = int(sz.z);
  void *d = nullptr;
  CUDA_RUNTIME(cudaMalloc(&d, sizeof(Q)));
  CUDA_RUNTIME(cudaMemcpy(d, &Q, sizeof(Q), cudaMemcpyHostToDevice));
  return d;
}
} // namespace

AstarothSim::AstarothSim(size_t x, size_t y, size_t z, const Options &opt) : opt_(opt) {
  if (opt_.fields < 1 || opt_.fields > kMaxFields) LOG_FATAL("AstarothSim supports 1.." << kMaxFields << " fields");
  dd_.reset(new DistributedDomain(x, y, z));
  if (!opt_.gpus.empty()) dd_->set_gpus(opt_.gpus);
  dd_->set_radius(kR);
  dd_->set_methods(opt_.methods);
  dd_->set_placement(opt_.placement);
  dd_->set_plan_file(opt_.planFile);
  for (int f = 0; f < opt_.fields; ++f) dd_->add_data(opt_.fp64 ? sizeof(double) : sizeof(float), "d" + std::to_string(f));
  dd_->realize();
  impl_.reset(new Impl);
  for (LocalDomain &ld : dd_->domains()) impl_->streams.emplace_back(ld.gpu());
  impl_->interior = dd_->get_interior();
  impl_->exterior = dd_->get_exterior();
  if (std::getenv("STENCIL_ASTAROTH_NO_TMA")) opt_.useTma = false;
  for (LocalDomain &ld : dd_->domains()) {
    DevGuard g(ld.gpu());
    cudaEvent_t e;
    CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    impl_->computeDone.push_back(e);
    impl_->computeRecorded.push_back(false);
  }
  if (opt_.autoSchedule && opt_.useTma && opt_.stencil == Stencil::Star19) opt_.overlap = false;
  impl_->freeRun = !opt_.overlap && opt_.freeRunning && dd_->exchange_device_only() && !std::getenv("STENCIL_ASTAROTH_HOST_SYNC");
  {
    /* whole-region TMA launches only, every subdomain its own x neighbour, rows that keep the second store aligned */
    const int n = opt_.fp64 ? TmaGeom<double>::N : TmaGeom<float>::N, tileX = 32 * n;
    bool ok = !opt_.overlap && opt_.fusedXFaces && opt_.useTma && opt_.stencil == Stencil::Star19 && dd_->subdomain_grid().x == 1 &&
              dd_->exchange_x_faces_all_local() && !std::getenv("STENCIL_ASTAROTH_NO_FUSED_X");
    for (LocalDomain &ld : dd_->domains()) ok = ok && ld.size().x >= tileX / 2 && ld.size().x >= 2 * kR + n && ld.size().y >= kTRows / 2;
    impl_->fusedX = ok;
  }
  setup_fused_faces();
  if (opt_.useTma && opt_.stencil == Stencil::Star19) {
    /* one tensor map per field and buffer: the padded buffer as (row pitch, rows, planes) */
    for (LocalDomain &ld : dd_->domains()) {
      DevGuard g(ld.gpu());
      const Dim3 raw = ld.raw_size();
      for (int parity = 0; parity < 2; ++parity) {
        std::vector<CUtensorMap> host;
        for (int f = 0; f < opt_.fields; ++f) {
          char *buf = static_cast<char *>(ld.slab()) + ld.buffer_start_offset(size_t(f), parity);
          const Dim3 pitch = ld.pitch(size_t(f));
          host.push_back(opt_.fp64 ? tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, pitch.x, raw.y, raw.z, TmaGeom<double>::boxX, TmaGeom<double>::boxY)
                                   : tma::make_plane_map(buf, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, pitch.x, raw.y, raw.z, TmaGeom<float>::boxX, TmaGeom<float>::boxY));
        }
        CUtensorMap *d = nullptr;
        CUDA_RUNTIME(cudaMalloc(&d, host.size() * sizeof(CUtensorMap)));
        CUDA_RUNTIME(cudaMemcpy(d, host.data(), host.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
        impl_->maps[parity].push_back(d);
      }
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, TmaGeom<double>::smemBytes));
      CUDA_RUNTIME(cudaFuncSetAttribute(astaroth_tma_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, TmaGeom<float>::smemBytes));
    }
  }
  init();
}

void AstarothSim::setup_fused_faces() {
  std::vector<LocalDomain> &doms = dd_->domains();
  impl_->push.resize(doms.size());
  bool ok = impl_->fusedX && impl_->freeRun && opt_.fusedYZFaces && !std::getenv("STENCIL_ASTAROTH_NO_FUSED_YZ");
  std::vector<std::vector<NeighborInfo>> all(doms.size()); // [di][f * kMaxFields + q]
  if (ok)
    for (size_t di = 0; di < doms.size() && ok; ++di) {
      all[di].resize(4 * kMaxFields);
      if (doms[di].size().y < 2 * kR || doms[di].size().z < 2 * kR || opt_.tileZ < kR) ok = false;
      for (int f = 0; f < 4 && ok; ++f)
        for (int q = 0; q < opt_.fields; ++q) {
          const NeighborInfo n = dd_->neighbor(di, kPushDir[f], size_t(q));
          all[di][size_t(f * kMaxFields + q)] = n;
          /* every face needs a neighbour we can store into; two processes on one GPU cannot hand-shake by spinning */
          if (!n.exists || !n.reachable || (n.sameDevice && !n.sameRank) || n.size.x != doms[di].size().x) ok = false;
        }
    }
  ok = dd_->comm().allreduce_sum(ok ? 0 : 1) == 0;
  impl_->fusedYZ = ok;
  if (!ok) return;
  CUDA_RUNTIME(cudaHostAlloc(&impl_->err, sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable));
  *impl_->err = 0;
  const int n = opt_.fp64 ? TmaGeom<double>::N : TmaGeom<float>::N, tileX = 32 * n;
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    DevGuard g(ld.gpu());
    Impl::Push &p = impl_->push[di];
    NeighborInfo nbs[4][kMaxFields];
    std::set<int> deps;
    p.finalWait.n = 0;
    for (int f = 0; f < 4; ++f) {
      for (int q = 0; q < opt_.fields; ++q) nbs[f][q] = all[di][size_t(f * kMaxFields + q)];
      const NeighborInfo &nb = nbs[f][0];
      p.faceMask |= 1u << f;
      if (!(nb.sameRank && nb.sameDevice)) p.flagMask |= 1u << f;
      if (nb.sameRank && nb.sameDevice && nb.dom != int(di)) deps.insert(nb.dom);
    }
    p.sameDevDeps.assign(deps.begin(), deps.end());
    /* CTAs of a whole-region launch that take part in the flag hand-shake (same classification as the kernel) */
    const Dim3 sz = ld.size();
    const int tilesX = int((sz.x + tileX - 1) / tileX), tilesY = int((sz.y + kTRows - 1) / kTRows);
    const int nChunks = int((sz.z + opt_.tileZ - 1) / opt_.tileZ);
    unsigned nBoundary = 0;
    for (int c = 0; c < nChunks; ++c)
      for (int ty = 0; ty < tilesY; ++ty) {
        const int z0 = c * opt_.tileZ, z1 = std::min<int>(z0 + opt_.tileZ, int(sz.z));
        if (tile_faces(ty, z0, z1, int(sz.y), int(sz.z), p.faceMask) & p.flagMask) nBoundary += unsigned(tilesX * opt_.fields);
      }
    CUDA_RUNTIME(cudaMalloc(&p.done, sizeof(unsigned)));
    CUDA_RUNTIME(cudaMemset(p.done, 0, sizeof(unsigned)));
    for (int parity = 0; parity < 2; ++parity)
      p.dev[parity] = opt_.fp64 ? build_push<double>(*dd_, di, parity, opt_.fields, nbs, p.faceMask, p.flagMask, nBoundary, p.done, impl_->err)
                                : build_push<float>(*dd_, di, parity, opt_.fields, nbs, p.faceMask, p.flagMask, nBoundary, p.done, impl_->err);
    for (int f = 0; f < 4; ++f)
      if ((p.flagMask >> f) & 1u) p.finalWait.flag[p.finalWait.n++] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(kPushDir[f]);
  }
}

AstarothSim::~AstarothSim() {
  if (dd_) sync();
  if (impl_ && dd_) {
    dd_->comm().barrier(); // nobody frees a slab that a neighbour's kernel may still be storing into
    for (size_t di = 0; di < impl_->push.size(); ++di) {
      DevGuard g(dd_->domains()[di].gpu());
      cudaFree(impl_->push[di].dev[0]);
      cudaFree(impl_->push[di].dev[1]);
      cudaFree(impl_->push[di].done);
    }
    if (impl_->err) cudaFreeHost(impl_->err);
  }
  if (impl_)
    for (cudaEvent_t e : impl_->computeDone) cudaEventDestroy(e);
  if (impl_ && dd_)
    for (int parity = 0; parity < 2; ++parity)
      for (size_t di = 0; di < impl_->maps[parity].size(); ++di) {
        DevGuard g(dd_->domains()[di].gpu());
        cudaFree(impl_->maps[parity][di]);
      }
  impl_.reset();
  dd_.reset();
}

void AstarothSim::init() {
  if (impl_->iter > 0) sync(); // neighbours may still be storing the last step's faces into these buffers
  for (size_t di = 0; di < dd_->domains().size(); ++di) {
    LocalDomain &ld = dd_->domains()[di];
    DevGuard g(ld.gpu());
    const Dim3 raw = ld.raw_size();
    const dim3 block(64, 4, 1), grid(unsigned((raw.x + 63) / 64), unsigned((raw.y + 3) / 4), unsigned(std::min<int64_t>(raw.z, 64)));
    for (int f = 0; f < opt_.fields; ++f) {
      const Dim3 pitch = ld.pitch(size_t(f));
      if (opt_.fp64)
        astaroth_init_kernel<double><<<grid, block, 0, impl_->streams[di]>>>(static_cast<double *>(ld.curr_data(size_t(f))), pitch.x, pitch.x * pitch.y,
                                                                              raw, ld.halo_origin(), opt_.period);
      else
        astaroth_init_kernel<float><<<grid, block, 0, impl_->streams[di]>>>(static_cast<float *>(ld.curr_data(size_t(f))), pitch.x, pitch.x * pitch.y,
                                                                             raw, ld.halo_origin(), opt_.period);
      CUDA_RUNTIME(cudaGetLastError());
      ++launches_;
    }
  }
  sync();
  impl_->xHalosFromKernel = false; // curr was just rewritten: the next exchange carries the x faces again
}

template <typename T> static void launch_region(AstarothSim::Stencil st, LocalDomain &ld, int fields, const Rect3 &reg, int tileZ, T dt, cudaStream_t stream) {
  if (reg.empty()) return;
  FieldParams<T> P;
  std::memset(&P, 0, sizeof(P));
  const Dim3 pitch = ld.pitch(0), org = ld.origin();
  const long long interiorOff = (long long)kR * (pitch.x * pitch.y + pitch.x + 1);
  for (int f = 0; f < fields; ++f) {
    P.src[f] = static_cast<const T *>(ld.curr_data(size_t(f))) + interiorOff;
    P.dst[f] = static_cast<T *>(ld.next_data(size_t(f))) + interiorOff;
  }
  P.pitchX = pitch.x;
  P.pitchXY = pitch.x * pitch.y;
  P.lox = int(reg.lo.x - org.x);
  P.loy = int(reg.lo.y - org.y);
  P.loz = int(reg.lo.z - org.z);
  P.hix = int(reg.hi.x - org.x);
  P.hiy = int(reg.hi.y - org.y);
  P.hiz = int(reg.hi.z - org.z);
  P.tileZ = tileZ;
  P.nChunks = (P.hiz - P.loz + tileZ - 1) / tileZ;
  P.dt = dt;
  constexpr int N = Vec<T>::n;
  const int xSpan = P.hix - (P.lox / N) * N;
  const dim3 block(kLanes, kRows, 1), grid(unsigned((xSpan + kLanes * N - 1) / (kLanes * N)), unsigned((P.hiy - P.loy + kRows - 1) / kRows),
                                           unsigned(P.nChunks * fields));
  if (st == AstarothSim::Stencil::Star19)
    astaroth_kernel<T, true><<<grid, block, 0, stream>>>(P);
  else
    astaroth_kernel<T, false><<<grid, block, 0, stream>>>(P);
  CUDA_RUNTIME(cudaGetLastError());
}

/* the TMA kernel on a region that is wide enough to fill its tiles */
template <typename T>
static void launch_region_tma(LocalDomain &ld, int fields, const CUtensorMap *maps, const Rect3 &reg, int tileZ, T dt, cudaStream_t stream,
                              bool wrapX = false, const void *push = nullptr, unsigned faceMask = 0, unsigned flagMask = 0, unsigned epoch = 0) {
  typedef TmaGeom<T> G;
  TmaParams<T> P;
  std::memset(&P, 0, sizeof(P));
  const Dim3 pitch = ld.pitch(0), org = ld.origin();
  const long long interiorOff = (long long)kR * (pitch.x * pitch.y + pitch.x + 1);
  for (int f = 0; f < fields; ++f) P.dst[f] = static_cast<T *>(ld.next_data(size_t(f))) + interiorOff;
  P.maps = maps;
  P.pitchX = pitch.x;
  P.pitchXY = pitch.x * pitch.y;
  P.lox = int(reg.lo.x - org.x);
  P.loy = int(reg.lo.y - org.y);
  P.loz = int(reg.lo.z - org.z);
  P.hix = int(reg.hi.x - org.x);
  P.hiy = int(reg.hi.y - org.y);
  P.hiz = int(reg.hi.z - org.z);
  P.tileZ = tileZ;
  P.tx0 = P.lox / G::tileX;
  P.ty0 = P.loy / kTRows;
  P.tilesX = (P.hix - 1) / G::tileX - P.tx0 + 1;
  P.tilesY = (P.hiy - 1) / kTRows - P.ty0 + 1;
  P.nChunks = (P.hiz - P.loz + tileZ - 1) / tileZ;
  P.x0pad = int(ld.lead_elems(0)) + kR;
  P.dt = dt;
  P.wrapX = wrapX ? int(ld.size().x) : 0;
  P.push = static_cast<const PushParams<T> *>(push);
  P.pushFaceMask = faceMask;
  P.pushFlagMask = flagMask;
  P.epoch = epoch;
  P.sy = int(ld.size().y);
  P.sz = int(ld.size().z);
  const dim3 grid(unsigned(P.tilesX * P.tilesY * P.nChunks), unsigned(fields), 1);
  astaroth_tma_kernel<T><<<grid, kTThreads, G::smemBytes, stream>>>(P);
  CUDA_RUNTIME(cudaGetLastError());
}

void AstarothSim::step() {
  STENCIL_RANGE("AstarothSim::step");
  std::vector<LocalDomain> &doms = dd_->domains();
  auto launch = [&](size_t di, const Rect3 &reg) {
    if (reg.empty()) return;
    DevGuard g(doms[di].gpu());
    const Dim3 ext = reg.extent();
    const int tileX = opt_.fp64 ? TmaGeom<double>::tileX : TmaGeom<float>::tileX;
    if (opt_.useTma && opt_.stencil == Stencil::Star19 && ext.x >= tileX / 2 && ext.y >= kTRows / 2) {
      const CUtensorMap *maps = impl_->maps[doms[di].parity()][di];
      const void *pushDev = impl_->fusedYZ ? impl_->push[di].dev[doms[di].parity()] : nullptr;
      const unsigned pushFaces = impl_->fusedYZ ? impl_->push[di].faceMask : 0u, pushFlags = impl_->fusedYZ ? impl_->push[di].flagMask : 0u;
      if (opt_.fp64)
        launch_region_tma<double>(doms[di], opt_.fields, maps, reg, opt_.tileZ, opt_.dt, impl_->streams[di], impl_->fusedX, pushDev, pushFaces,
                                  pushFlags, impl_->iter);
      else
        launch_region_tma<float>(doms[di], opt_.fields, maps, reg, opt_.tileZ, float(opt_.dt), impl_->streams[di], impl_->fusedX, pushDev, pushFaces,
                                 pushFlags, impl_->iter);
      ++launches_;
      return;
    }
    if (opt_.fp64)
      launch_region<double>(opt_.stencil, doms[di], opt_.fields, reg, opt_.tileZ, opt_.dt, impl_->streams[di]);
    else
      launch_region<float>(opt_.stencil, doms[di], opt_.fields, reg, opt_.tileZ, float(opt_.dt), impl_->streams[di]);
    ++launches_;
  };
  /* x halos of `curr`: written by the previous step's kernel if there was one, else by this exchange */
  dd_->exchange_skip_local_x_faces(impl_->fusedX && impl_->xHalosFromKernel);
  const bool steady = impl_->fusedYZ && impl_->xHalosFromKernel; // every halo of `curr` was written by the previous kernels
  impl_->xHalosFromKernel = impl_->fusedX;
  ++impl_->iter;
  if (impl_->freeRun) {
    /* exchange_async() on the exchange streams, then one whole-region launch per subdomain ordered after it by an
       event; the next exchange is ordered after these launches the same way. The host only enqueues.
       With fused faces there is no exchange at all after the first step: kernels of different GPUs hand-shake through
       flags, kernels of subdomains that share a GPU are ordered by events. */
    if (!steady) {
      for (size_t di = 0; di < doms.size(); ++di)
        if (impl_->computeRecorded[di]) dd_->exchange_after(di, impl_->computeDone[di]);
      dd_->exchange_async();
    }
    for (size_t di = 0; di < doms.size(); ++di) {
      DevGuard g(doms[di].gpu());
      if (!steady) {
        CUDA_RUNTIME(cudaStreamWaitEvent(impl_->streams[di], dd_->exchange_done_event(di), 0));
      } else {
        for (int dep : impl_->push[di].sameDevDeps)
          if (impl_->computeRecorded[size_t(dep)]) CUDA_RUNTIME(cudaStreamWaitEvent(impl_->streams[di], impl_->computeDone[size_t(dep)], 0));
      }
    }
    for (size_t di = 0; di < doms.size(); ++di) {
      DevGuard g(doms[di].gpu());
      launch(di, doms[di].get_compute_region());
      CUDA_RUNTIME(cudaEventRecord(impl_->computeDone[di], impl_->streams[di]));
      impl_->computeRecorded[di] = true;
    }
    dd_->swap();
    return;
  }
  if (opt_.overlap)
    for (size_t di = 0; di < doms.size(); ++di) launch(di, impl_->interior[di]);
  const double t0 = Comm::wtime();
  dd_->exchange();
  impl_->exchangeSeconds += Comm::wtime() - t0;
  for (size_t di = 0; di < doms.size(); ++di) {
    if (opt_.overlap)
      for (const Rect3 &r : impl_->exterior[di]) launch(di, r);
    else
      launch(di, doms[di].get_compute_region());
  }
  sync();
  dd_->swap();
}

void AstarothSim::sync() {
  for (size_t di = 0; di < dd_->domains().size(); ++di) {
    DevGuard g(dd_->domains()[di].gpu());
    /* fused faces: my halos of the last iteration are complete only when every neighbour GPU raised its last flag */
    if (impl_->fusedYZ && impl_->iter > 0 && impl_->xHalosFromKernel)
      launch_wait_flags(impl_->push[di].finalWait, impl_->iter, impl_->err, 5000000000ull, impl_->streams[di]);
    CUDA_RUNTIME(cudaStreamSynchronize(impl_->streams[di]));
  }
  dd_->exchange_wait(); // no-op unless an asynchronous exchange is still in flight
  if (impl_->err && *impl_->err) LOG_FATAL("AstarothSim: timed out waiting for a neighbour GPU (code " << *impl_->err << ", iteration " << impl_->iter << ")");
}

double AstarothSim::run(int iters, double *exchangeSeconds) {
  sync();
  dd_->comm().barrier();
  impl_->exchangeSeconds = 0;
  const double t0 = Comm::wtime();
  for (int i = 0; i < iters; ++i) step();
  if (impl_->freeRun) sync();
  const double t = dd_->comm().allreduce_max(Comm::wtime() - t0);
  if (impl_->freeRun) {
    /* the loop never stops for the exchange, so its share is reported as the cost of the same exchange run alone */
    dd_->comm().barrier();
    const double e0 = Comm::wtime();
    dd_->exchange();
    impl_->exchangeSeconds = (Comm::wtime() - e0) * iters;
  }
  if (exchangeSeconds) *exchangeSeconds = dd_->comm().allreduce_max(impl_->exchangeSeconds);
  return t;
}

std::vector<double> AstarothSim::gather(int field) {
  sync();
  const Dim3 g = dd_->size();
  std::vector<double> out(size_t(g.flatten()), std::numeric_limits<double>::quiet_NaN());
  for (LocalDomain &ld : dd_->domains()) {
    const std::vector<unsigned char> b = ld.interior_to_host(size_t(field));
    const Dim3 sz = ld.size(), o = ld.origin();
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x, ++i) {
          const double v = opt_.fp64 ? reinterpret_cast<const double *>(b.data())[i] : double(reinterpret_cast<const float *>(b.data())[i]);
          out[size_t(((o.z + z) * g.y + o.y + y) * g.x + o.x + x)] = v;
        }
  }
  return out;
}

double astaroth_init_value(const Dim3 &p, double period, bool fp64) {
  const double w = 2 * 3.14159 / period;
  const double v = std::sin(w * double(p.x) + w * double(p.y) + w * double(p.z));
  return fp64 ? v : double(float(v));
}

template <typename T> static void host_step(std::vector<double> &next, const std::vector<double> &curr, const Dim3 &size, AstarothSim::Stencil st, T dt) {
  const int64_t X = size.x, Y = size.y, Z = size.z;
  next.resize(curr.size());
  auto at = [&](int64_t x, int64_t y, int64_t z) { return T(curr[size_t((((z + 3 * Z) % Z) * Y + (y + 3 * Y) % Y) * X + (x + 3 * X) % X)]); };
  for (int64_t z = 0; z < Z; ++z)
    for (int64_t y = 0; y < Y; ++y)
      for (int64_t x = 0; x < X; ++x) {
        T v;
        if (st == AstarothSim::Stencil::Star19) {
          T xm[3], xp[3], ym[3], yp[3], zm[3], zp[3];
          for (int k = 0; k < 3; ++k) {
            xm[k] = at(x - k - 1, y, z);
            xp[k] = at(x + k + 1, y, z);
            ym[k] = at(x, y - k - 1, z);
            yp[k] = at(x, y + k + 1, z);
            zm[k] = at(x, y, z - k - 1);
            zp[k] = at(x, y, z + k + 1);
          }
          v = star19<T>(at(x, y, z), xm, xp, ym, yp, zm, zp, dt);
        } else {
          v = average7<T>(at(x - 1, y, z), at(x, y - 1, z), at(x, y, z - 1), at(x + 1, y, z), at(x, y + 1, z), at(x, y, z + 1));
        }
        next[size_t((z * Y + y) * X + x)] = double(v);
      }
}

void astaroth_host_step(std::vector<double> &next, const std::vector<double> &curr, const Dim3 &size, AstarothSim::Stencil st, bool fp64, double dt) {
  if (fp64)
    host_step<double>(next, curr, size, st, dt);
  else
    host_step<float>(next, curr, size, st, float(dt));
}

} // namespace stencil
