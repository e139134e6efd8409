/* Jacobi3D engine and its sm_100a kernels (N2 jacobi7 fused exterior, N3 jacobi7 interior, N9 init).
   Replaces reference kernels K8/K9 (bin/jacobi3d.cu:18-87). See include/stencil/jacobi.hpp for the modes.

   Two kernel variants compute the same thing (Options::useTma picks; TMA is the default):

   jacobi7_tma_kernel — the Blackwell path. A tile is 128 x 16 cells in xy, marched along z. Thread 0 streams the
   (128+8) x (16+2) input planes of the tile into a 6-deep shared-memory ring with cp.async.bulk.tensor.3d (TMA;
   out-of-range parts of edge tiles are zero-filled by the hardware) and arms one mbarrier per stage; the eight warps
   wait on the stage's mbarrier, read their centre / +-y rows as 16-byte LDS, take +-x from the neighbouring lane, and
   hand the stage back through a second mbarrier. The centre column of the last two planes and the in-plane partial
   sum live in registers, so every plane is read from shared memory exactly once. Bytes in flight are set by the ring
   (6 x 9.8 KB per CTA, 2 CTAs per SM), not by registers, which is what a latency-bound streaming stencil needs to
   reach HBM speed.

   jacobi7_kernel — plain LDG version kept for comparison. A CTA is 32 lanes x 8 rows; a lane owns 4 consecutive x cells (one 16-byte vector), so a tile is
   128 x 8 cells in xy and is marched along z for tileZ planes with the centre column kept in registers
   (below / cur / above rotate). Per plane a thread issues three 16-byte loads (above, +y row, -y row); the x
   neighbours come from the adjacent lanes by shuffle, only lane 0 / lane 31 load one scalar each. Rows start on
   128-byte boundaries (LocalDomain layout), so every vector access is aligned and a warp reads whole cache lines.
   HBM traffic is one read + one write of the field per iteration; the +-y / +-z re-reads are served by L1/L2.

   Fusion. Tiles that own cells on a subdomain face additionally put those cells into the neighbour's `next` halo
   (peer pointer: NVLink store, or a plain local store when the axis wraps onto the same subdomain) — 16-byte vectors
   for y and z faces, one whole 32-byte sector per row (halo cell + row padding) for x faces. In the TMA kernel this
   is an epilogue of the tile: the plane loop is the plain kernel's loop, and the face cells are read back from `next`
   (they are still in L2) once the loop is done; the LDG variant stores them from registers as it goes. Tiles that
   stored into ANOTHER GPU finish with a GPU-scope fence + a counter increment; the last of them issues one
   system-scope fence and writes the iteration number to every such neighbour's flag. At the start of the next
   iteration only those tiles look at the flags (ld.acquire.sys): they are the only ones that read halo cells written
   by another GPU, and their neighbours' flags were raised well before the previous kernel ended, so in steady state
   nobody spins. Same-GPU neighbours are ordered by the stream. */
#include "stencil/detail/tma.cuh"
#include "stencil/nvtx.hpp"
#include "stencil/jacobi.hpp"

#include <algorithm>
#include <cstring>
#include <set>

#include <cuda.h>

#include "stencil/copy_plan.hpp"
#include "stencil/detail/launch_pool.hpp"
#include "stencil/reduce.hpp"

namespace stencil {

namespace {

constexpr int kLanes = 32, kRows = 8, kThreads = kLanes * kRows, kVec = 4, kTileX = kLanes * kVec;
/* TMA variant geometry. Two tile shapes share one kernel, picked per tile (uniform per CTA):
     wide   128 cells x 16 rows: a warp is one pair of rows, a lane owns 4 cells of each;
     narrow  32 cells x 64 rows: a warp is four pairs of rows, 8 lanes per row. Same 2048 cells per plane, same loop.
   Narrow tiles cover what is left of a row after the last full 128-cell tile: the weak-scaling rule produces rows of
   645 and 813 cells, and a wide tile on the remaining 5 / 45 cells costs as much as one on 128 (the kernel is as much
   issue bound as DRAM bound). 645 -> 5 wide + 1 narrow column = 5.25 tile columns instead of 6. */
constexpr int kTmaRows = 16, kTmaStages = 6;
constexpr int kTmaConsumerWarps = kTmaRows / 2, kTmaThreads = kTmaConsumerWarps * 32;
constexpr int kAutoTileZ = 48; // layer depth chosen for large subdomains when Options::tileZ is 0 (measured: profiles/README.md)
constexpr int kMaxTileZ = 64; // bound on Options::tileZ for the TMA variant (keeps enough tiles in flight on small domains)
constexpr int kTmaLag = 2; // the producer refills the slot that was consumed kTmaLag planes ago
/* Subdomains that wrap onto themselves in x (always, with the default partition) do not keep x halo cells at all in
   fused mode: a tile at the left / right end of the rows loads the four cells at the OTHER end of its rows with a
   second, tiny TMA box per plane (4 cells x tile rows) and takes its x = -1 / x = sx neighbour from there. The far-end
   cells are being read by the tile at that end at about the same time, so this is an L2 hit; it replaces one 32-byte
   sector store per row and its read-back. */
constexpr int kSideBoxX = 4;
template <int LW> struct Geo { // LW = lanes per row: 32 (wide) or 8 (narrow)
  static constexpr int kSub = 32 / LW;         // row pairs per warp
  static constexpr int kW = LW * kVec;         // tile width in cells
  static constexpr int kR = kTmaRows * kSub;   // tile rows
  static constexpr int kBoxX = kW + 8, kBoxY = kR + 2;
  static constexpr int kPlaneBytes = kBoxX * kBoxY * 4;
  static constexpr int kStageBytes = (kPlaneBytes + 127) / 128 * 128;
  static constexpr int kSideBytes = kSideBoxX * kR * 4; // per side and stage
  static constexpr int kSideOffset = kTmaStages * kStageBytes;
  static constexpr int kBarOffset = kSideOffset + 2 * kTmaStages * kSideBytes;
  static constexpr int kSmemBytes = kBarOffset + 2 * kTmaStages * 8;
};
using Wide = Geo<32>;
using Narrow = Geo<8>;
constexpr int kNarrowMaxRemainder = 96; // a remainder wider than this is cheaper as one wide tile than as four narrow ones
constexpr int kTmaSmemBytes = Narrow::kSmemBytes > Wide::kSmemBytes ? Narrow::kSmemBytes : Wide::kSmemBytes;
constexpr size_t kDynBytes = Jacobi3D::kParamBytes;
constexpr int kFusedFlagBase = 64; // word offset of the fused-iteration flags inside a control block

enum Face { XM = 0, XP = 1, YM = 2, YP = 3, ZM = 4, ZP = 5 };

struct Tile {
  int x0, y0, z0;       // first cell / first row / first plane of the tile (compute coordinates)
  unsigned short nz;    // planes the tile marches (layers differ in depth: see z_layers)
  unsigned char faces;  // bit f set when the tile owns cells on face f that must be pushed
  unsigned char narrow; // TMA variant: 1 = 32 x 64 tile, 0 = 128 x 16
};
static_assert(sizeof(Tile) == 16, "one 16-byte load per CTA");

struct JacobiParams {
  const float *src; // interior origin: compute cell (0,0,0) of curr
  float *dst;       // interior origin of next
  long long pitchX, pitchXY; // elements: row pitch, plane pitch
  int sx, sy, sz;            // compute-region size
  int lox, loy, loz, hix, hiy, hiz; // cells to update (compute coordinates, half open)
  int wrapLoad;       // fused mode, subdomain is its own +-x neighbour: x neighbours across the wrap come from side TMA boxes
  unsigned debugSkip; // diagnostics only (STENCIL_JACOBI_DEBUG_SKIP): 1 no signal, 2 no x push, 4 no y/z push,
                      // 8 wide tiles fetch 16 rows per plane instead of 18 (WRONG results: timing bound for halo-row sharing)
  int tmaX0; // index of compute cell x=0 inside a padded row (TMA coordinates count from the row start)
  /* Dirichlet spheres in global coordinates */
  long long ox, oy, oz;   // global coordinate of compute cell (0,0,0)
  long long hcx, ccx, cy, cz, r2; // hot/cold centre x, shared centre y/z, (R+1)^2
  int sphere32;           // every squared distance of the sphere tests fits 32 bits (domains up to 30000 cells per axis)
  const Tile *tiles;
  const float *dyn; // device-resident run-time parameters: [0] hot temperature, [1] cold temperature
  /* fused push */
  float *nb[6];            // neighbour's next buffer: address of ITS halo cell that receives my face cell (0,0) ...
  long long nbPitchX[6], nbPitchXY[6];
  unsigned *nbFlag[6];     // neighbour flag to raise (nullptr: ordered by the stream / same kernel)
  const unsigned *myFlag[6]; // my flag the neighbour raises (nullptr: no wait)
  unsigned flagMask;       // faces whose neighbour is synchronised through a flag (another GPU); 0 on a single GPU
  unsigned epoch;          // iteration number being computed (1-based)
  unsigned nFace[6];       // per face: number of tiles that push through it to another GPU
  unsigned *done;          // six per-face counters of finished tiles
  unsigned *err;           // host-mapped error word
  unsigned long long timeoutNs;
};

/* Plain (coherent) loads on purpose: halo cells are written by OTHER GPUs while this kernel may already be running,
   and only become readable after the acquire on the neighbour's flag. ld.global.nc is outside that guarantee. */
__device__ __forceinline__ float4 ldg4(const float *p) { return *reinterpret_cast<const float4 *>(p); }
/* relaxed: always issued right after one system-scope fence (a release store per flag would be a fence each) */
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

/* x / 6 correctly rounded in three instructions (Markstein: q = RN(x*y), r = x - 6q exactly by FMA, RN(q + r*y) is the
   correctly rounded quotient when y = RN(1/6)); bit-identical to IEEE division for the normal range, which is all a
   field bounded by the sphere temperatures produces. The compiler's generic division sequence costs ~4x more issue
   slots, and issue slots — not HBM — are what this kernel would otherwise run out of first. */
__device__ __forceinline__ float div6(float x) {
  const float y = 0x1.555556p-3f; // RN(1/6)
  const float q = x * y;
  const float r = __fmaf_rn(-6.0f, q, x);
  return __fmaf_rn(r, y, q);
}

/* one output value, in the reference's summation order, with the sphere override */
__device__ __forceinline__ float relax(float px, float mx, float py, float my, float pz, float mz) {
  float v = 0;
  v += px;
  v += mx;
  v += py;
  v += my;
  v += pz;
  v += mz;
  return div6(v);
}

__device__ __forceinline__ float sphere_or(const JacobiParams &P, long long gx, long long gy, long long gz, float v) {
  const long long dy = gy - P.cy, dz = gz - P.cz;
  const long long base = dy * dy + dz * dz;
  const long long dh = gx - P.hcx, dc = gx - P.ccx;
  /* trunc(sqrt(float(d2))) <= R  <=>  d2 < (R+1)^2 for integers below 2^24 */
  if (base + dh * dh < P.r2) return P.dyn[0];
  if (base + dc * dc < P.r2) return P.dyn[1];
  return v;
}

/* Tiles on a flagged face read halo cells another GPU writes: wait until that neighbour has delivered the previous
   iteration through THAT face (threads 0..5, one per face; the caller follows with a CTA barrier). A tile waits only
   on the faces it touches, so a neighbour that is late on one face holds up only the tiles along that face. */
__device__ __forceinline__ void wait_faces(const JacobiParams &P, unsigned faces) {
  if (threadIdx.x < 6 && ((faces >> threadIdx.x) & 1u)) {
    const unsigned *f = P.myFlag[threadIdx.x];
    const unsigned long long t0 = globaltimer_ns();
    while (int(ld_acquire_sys(f) - (P.epoch - 1)) < 0) {
      __nanosleep(64);
      if (globaltimer_ns() - t0 > P.timeoutNs) {
        atomicCAS(P.err, 0u, 100u + threadIdx.x);
        break;
      }
    }
  }
}

/* Tiles that stored into ANOTHER GPU: count this CTA as done on each of those faces; whoever completes a face's count
   raises that neighbour's flag — per face, so a face whose tiles are scheduled early (see make_tiles) is announced
   early and the neighbour never finds it missing when its next kernel starts.
   Ordering chain: every thread's (peer) stores -> CTA barrier -> thread 0: fence at GPU scope + counter increments ->
   the CTA that sees a full count: system-scope fence -> relaxed stores of the flags. Causality is transitive
   across the scopes, so only the final fence has to be system-wide; a system fence per boundary CTA measured
   ~20 us per 512^3 iteration on B200. */
__device__ __forceinline__ void signal_faces(const JacobiParams &P, unsigned faces) {
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned ready = 0;
#pragma unroll
    for (int f = 0; f < 6; ++f)
      if ((faces >> f) & 1u) {
        if (atomicAdd(P.done + f, 1u) == P.nFace[f] - 1) {
          P.done[f] = 0;
          ready |= 1u << f;
        }
      }
    if (ready) {
      __threadfence_system();
#pragma unroll
      for (int f = 0; f < 6; ++f)
        if ((ready >> f) & 1u) st_flag_sys(P.nbFlag[f], P.epoch);
    }
  }
}

template <bool FUSED> __global__ void __launch_bounds__(kThreads) jacobi7_kernel(const __grid_constant__ JacobiParams P) {
  const Tile t = P.tiles[blockIdx.x];
  const int lane = threadIdx.x & (kLanes - 1), row = threadIdx.x / kLanes;

  if (FUSED && (t.faces & P.flagMask)) {
    wait_faces(P, t.faces & P.flagMask);
    __syncthreads();
  }

  const int x0 = t.x0 + lane * kVec;
  const int y = t.y0 + row;
  const int z0 = max(t.z0, P.loz), z1 = min(t.z0 + int(t.nz), P.hiz);
  /* a warp is one row, so the row test is warp-uniform; the x test is per lane (partial tiles) */
  const bool rowActive = y >= P.loy && y < P.hiy;
  const bool laneActive = x0 < P.hix && x0 + kVec > P.lox;

  /* does this tile come near either sphere? (uniform per CTA) */
  bool nearSphere;
  {
    const long long gy0 = P.oy + t.y0, gy1 = gy0 + kRows - 1;
    const long long gz0 = P.oz + z0, gz1 = P.oz + z1 - 1;
    const long long gx0 = P.ox + t.x0, gx1 = gx0 + kTileX - 1;
    auto gap = [](long long a0, long long a1, long long c) { return c < a0 ? a0 - c : (c > a1 ? c - a1 : 0); };
    const long long dy = gap(gy0, gy1, P.cy), dz = gap(gz0, gz1, P.cz);
    const long long dh = gap(gx0, gx1, P.hcx), dc = gap(gx0, gx1, P.ccx);
    nearSphere = (dy * dy + dz * dz + dh * dh < P.r2) || (dy * dy + dz * dz + dc * dc < P.r2);
  }

  if (rowActive && z0 < z1) {
    const float *s = P.src + (long long)y * P.pitchX + x0;
    float *d = P.dst + (long long)y * P.pitchX + x0;
    const bool full = x0 >= P.lox && x0 + kVec <= P.hix;
    /* the x neighbours come from the adjacent lane unless that lane is outside the tile or the region */
    const bool needW = lane == 0 || x0 <= P.lox, needE = lane == kLanes - 1 || x0 + kVec >= P.hix;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 below = laneActive ? ldg4(s + (long long)(z0 - 1) * P.pitchXY) : zero4;
    float4 cur = laneActive ? ldg4(s + (long long)z0 * P.pitchXY) : zero4;
#pragma unroll 2
    for (int z = z0; z < z1; ++z) {
      const float *sp = s + (long long)z * P.pitchXY;
      /* every lane of the warp takes part in the shuffles; everything else is per active lane */
      float west = __shfl_up_sync(0xffffffffu, cur.w, 1);
      float east = __shfl_down_sync(0xffffffffu, cur.x, 1);
      if (!laneActive) continue;
      const float4 above = ldg4(sp + P.pitchXY);
      const float4 north = ldg4(sp + P.pitchX);
      const float4 south = ldg4(sp - P.pitchX);
      if (needW) west = sp[-1];
      if (needE) east = sp[kVec];
      float4 o;
      o.x = relax(cur.y, west, north.x, south.x, above.x, below.x);
      o.y = relax(cur.z, cur.x, north.y, south.y, above.y, below.y);
      o.z = relax(cur.w, cur.y, north.z, south.z, above.z, below.z);
      o.w = relax(east, cur.z, north.w, south.w, above.w, below.w);
      if (nearSphere) {
        const long long gx = P.ox + x0, gy = P.oy + y, gz = P.oz + z;
        o.x = sphere_or(P, gx, gy, gz, o.x);
        o.y = sphere_or(P, gx + 1, gy, gz, o.y);
        o.z = sphere_or(P, gx + 2, gy, gz, o.z);
        o.w = sphere_or(P, gx + 3, gy, gz, o.w);
      }
      float *dp = d + (long long)z * P.pitchXY;
      const float ov[4] = {o.x, o.y, o.z, o.w};
      if (full) {
        *reinterpret_cast<float4 *>(dp) = o;
      } else {
#pragma unroll
        for (int k = 0; k < kVec; ++k)
          if (x0 + k >= P.lox && x0 + k < P.hix) dp[k] = ov[k];
      }
      if (FUSED && t.faces) {
        /* neighbour halo addresses are pre-offset so that (y, z, x) of MY compute coordinates index them directly */
        if ((t.faces & (1u << YM)) && y == 0) {
          float *q = P.nb[YM] + (long long)z * P.nbPitchXY[YM] + x0;
          if (full) *reinterpret_cast<float4 *>(q) = o;
          else
            for (int k = 0; k < kVec; ++k)
              if (x0 + k < P.hix) q[k] = ov[k];
        }
        if ((t.faces & (1u << YP)) && y == P.sy - 1) {
          float *q = P.nb[YP] + (long long)z * P.nbPitchXY[YP] + x0;
          if (full) *reinterpret_cast<float4 *>(q) = o;
          else
            for (int k = 0; k < kVec; ++k)
              if (x0 + k < P.hix) q[k] = ov[k];
        }
        if ((t.faces & (1u << ZM)) && z == 0) {
          float *q = P.nb[ZM] + (long long)y * P.nbPitchX[ZM] + x0;
          if (full) *reinterpret_cast<float4 *>(q) = o;
          else
            for (int k = 0; k < kVec; ++k)
              if (x0 + k < P.hix) q[k] = ov[k];
        }
        if ((t.faces & (1u << ZP)) && z == P.sz - 1) {
          float *q = P.nb[ZP] + (long long)y * P.nbPitchX[ZP] + x0;
          if (full) *reinterpret_cast<float4 *>(q) = o;
          else
            for (int k = 0; k < kVec; ++k)
              if (x0 + k < P.hix) q[k] = ov[k];
        }
        if ((t.faces & (1u << XM)) && x0 == 0) P.nb[XM][(long long)z * P.nbPitchXY[XM] + (long long)y * P.nbPitchX[XM]] = o.x;
        if ((t.faces & (1u << XP)) && x0 <= P.sx - 1 && P.sx - 1 < x0 + kVec)
        {
          const int k = P.sx - 1 - x0; // which of my four cells is the last one of the row
          P.nb[XP][(long long)z * P.nbPitchXY[XP] + (long long)y * P.nbPitchX[XP]] = k == 0 ? o.x : (k == 1 ? o.y : (k == 2 ? o.z : o.w));
        }
      }
      below = cur;
      cur = above;
    }
  }

  if (FUSED && (t.faces & P.flagMask)) {
    __syncthreads();
    signal_faces(P, t.faces & P.flagMask);
  }
}


/* mbarrier / TMA primitives: stencil/detail/tma.cuh */
using tma::mbar_arrive;
using tma::mbar_arrive_expect_tx;
using tma::mbar_init;
using tma::mbar_wait;
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, int cx, int cy, int cz, unsigned long long *bar) {
  tma::load_3d(dst, map, cx, cy, cz, bar);
}

/* TMA kernel, boundary tiles, after the plane loop: copy the y / z face rows this tile has just written to `next`
   (still in L2) into the neighbours' halos. Whole CTA, one warp per row, 16-byte vectors. Keeping this out of the plane
   loop leaves the loop free of face tests and calls. tw x tr = the tile's shape. */
__device__ __noinline__ void push_tile_yz(const JacobiParams &P, unsigned faces, int xlo, int ylo, int tw, int tr, int z0, int z1) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = kTmaThreads / 32;
  const int xhi = min(xlo + tw, P.sx), yhi = min(ylo + tr, P.sy);
  for (int f = YM; f <= ZP; ++f) {
    if (!((faces >> f) & 1u)) continue;
    const bool yface = f == YM || f == YP;
    const int fixed = (f == YM || f == ZM) ? 0 : (yface ? P.sy - 1 : P.sz - 1);
    /* y face: one row per plane of the tile; z face: the rows of the tile in one plane */
    const int rows = yface ? z1 - z0 : yhi - ylo;
    for (int r = warp; r < rows; r += nwarps) {
      const int y = yface ? fixed : ylo + r, z = yface ? z0 + r : fixed;
      const float *src = P.dst + (long long)z * P.pitchXY + (long long)y * P.pitchX;
      float *dst = yface ? P.nb[f] + (long long)z * P.nbPitchXY[f] : P.nb[f] + (long long)y * P.nbPitchX[f];
      for (int x = xlo + lane * kVec; x < xhi; x += 32 * kVec) {
        if (x + kVec <= xhi) {
          *reinterpret_cast<float4 *>(dst + x) = __ldcg(reinterpret_cast<const float4 *>(src + x));
        } else {
          for (int i = x; i < xhi; ++i) dst[i] = __ldcg(src + i);
        }
      }
    }
  }
}

/* ... and the x faces (subdomains whose x neighbour is not themselves): one cell per row of the tile, so one cell per
   thread. The value is read back from `next` (L2) and goes into the neighbour's halo cell — as a whole 32-byte sector
   where the halo cell shares its sector with row padding only (LocalDomain layout), because a 4-byte store would make
   the receiver's memory system read-modify-write the sector. */
__device__ __noinline__ void push_tile_x(const JacobiParams &P, unsigned faces, int ylo, int tr, int z0, int z1) {
  const int n = (z1 - z0) * tr;
  for (int f = XM; f <= XP; ++f) {
    if (!((faces >> f) & 1u)) continue;
    const int x = f == XM ? 0 : P.sx - 1;
    /* start of the sector holding the destination halo cell: my x=0 lands in the neighbour's +x halo, the first cell
       after its interior (sector = halo + 7 padding cells); my last x lands in its -x halo, the last cell before its
       128-byte aligned interior (sector = 7 padding cells + halo) */
    const long long sectorShift = f == XM ? 0 : -7;
    const bool wide = ((reinterpret_cast<unsigned long long>(P.nb[f] + sectorShift) | (unsigned long long)(P.nbPitchX[f] * 4) |
                        (unsigned long long)(P.nbPitchXY[f] * 4)) & 31ull) == 0;
    for (int i = threadIdx.x; i < n; i += kTmaThreads) {
      const int z = z0 + i / tr, y = ylo + i % tr;
      if (y >= P.sy) continue;
      const float v = __ldcg(P.dst + (long long)z * P.pitchXY + (long long)y * P.pitchX + x);
      float *q = P.nb[f] + (long long)z * P.nbPitchXY[f] + (long long)y * P.nbPitchX[f];
      if (wide) {
        const float4 vv = make_float4(v, v, v, v);
        reinterpret_cast<float4 *>(q + sectorShift)[0] = vv;
        reinterpret_cast<float4 *>(q + sectorShift)[1] = vv;
      } else { // neighbour's x extent is not a multiple of 8 cells: interior cells share the sector
        *q = v;
      }
    }
  }
}

__device__ __forceinline__ bool tile_near_sphere(const JacobiParams &P, const Tile &t, int tw, int tr, int z0, int z1) {
  const long long gy0 = P.oy + t.y0, gy1 = gy0 + tr - 1;
  const long long gz0 = P.oz + z0, gz1 = P.oz + z1 - 1;
  const long long gx0 = P.ox + t.x0, gx1 = gx0 + tw - 1;
  auto gap = [](long long a0, long long a1, long long c) { return c < a0 ? a0 - c : (c > a1 ? c - a1 : 0); };
  const long long dy = gap(gy0, gy1, P.cy), dz = gap(gz0, gz1, P.cz);
  const long long dh = gap(gx0, gx1, P.hcx), dc = gap(gx0, gx1, P.ccx);
  return (dy * dy + dz * dz + dh * dh < P.r2) || (dy * dy + dz * dz + dc * dc < P.r2);
}

/* rare paths of the plane loop, kept out of line so they cost the hot path nothing but a branch */
__device__ __noinline__ float4 apply_spheres(const JacobiParams &P, int x0, int y, int z, float4 o) {
  /* by value in and out: passing the result array by reference would pin it in local memory for every plane */
  const long long gx = P.ox + x0, gy = P.oy + y, gz = P.oz + z;
  o.x = sphere_or(P, gx, gy, gz, o.x);
  o.y = sphere_or(P, gx + 1, gy, gz, o.y);
  o.z = sphere_or(P, gx + 2, gy, gz, o.z);
  o.w = sphere_or(P, gx + 3, gy, gz, o.w);
  return o;
}

__device__ __noinline__ void store_partial(const JacobiParams &P, float *dp, int x0, float4 o) {
  const float ov[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
  for (int k = 0; k < kVec; ++k)
    if (x0 + k >= P.lox && x0 + k < P.hix) dp[k] = ov[k];
}

/* one tile of the TMA kernel; G = Wide or Narrow (uniform per CTA) */
template <bool FUSED, class G>
__device__ __forceinline__ void tma_tile(unsigned char *smem, const CUtensorMap *tmap, const CUtensorMap *tmapSide, const JacobiParams &P, const Tile &t) {
  constexpr int LW = 32 / G::kSub;
  unsigned long long *full = reinterpret_cast<unsigned long long *>(smem + G::kBarOffset);
  unsigned long long *empty = full + kTmaStages;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int xl = lane & (LW - 1), sub = lane / LW; // position in the row, which of the warp's row pairs

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kTmaConsumerWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  /* Programmatic dependent launch (Jacobi3D::Options::pdl): this grid may have been started while the previous
     iteration's last CTAs were still running — its CTAs take the slots those free, set up their barriers and fetch their
     table entry meanwhile — and everything that touches the fields waits here for that grid to complete. A no-op in a
     launch without the attribute. */
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (FUSED && (t.faces & P.flagMask)) wait_faces(P, t.faces & P.flagMask);
  __syncthreads();
  /* The halo cells this tile is about to fetch were written by another GPU with ordinary (generic-proxy) stores and
     made visible by the acquire above; the fetch is a TMA (async-proxy) read issued by thread 0, so that thread orders
     the two proxies explicitly. */
  if (FUSED && (t.faces & P.flagMask) && threadIdx.x == 0) asm volatile("fence.proxy.async.global;" ::: "memory");

  const int z0 = max(t.z0, P.loz), z1 = min(t.z0 + int(t.nz), P.hiz);
  const int nplanes = z1 > z0 ? z1 - z0 + 2 : 0; // input planes z0-1 .. z1

  /* There is no producer warp: a ninth warp would cost a tenth warp's worth of registers (warps are allocated in
     pairs) and with it a resident CTA. Thread 0 issues the TMA loads itself — a prologue of kTmaStages - kTmaLag
     planes, then one refill per iteration into the slot every warp released kTmaLag planes ago, so its wait on the
     `empty` barrier is almost always already satisfied. */
  const int cx = P.tmaX0 + t.x0 - 4; // 4 cells of left margin keep the centre 16-byte aligned in smem
  const int cy = t.y0;               // raw row of compute row (y0 - 1)
  /* x wrap by load: does this tile hold the first / the last cell of the rows? */
  const bool sideL = FUSED && P.wrapLoad && t.x0 == 0;
  const bool sideR = FUSED && P.wrapLoad && t.x0 + G::kW >= P.sx;
  auto issue = [&](int j) {          // load input plane j (compute plane z0 - 1 + j, raw plane z0 + j)
    const int sj = j % kTmaStages;
    if (j >= kTmaStages) mbar_wait(&empty[sj], ((j / kTmaStages) - 1) & 1);
    const int planeBytes = (LW == 32 && (P.debugSkip & 8u)) ? G::kBoxX * (G::kBoxY - 2) * 4 : G::kPlaneBytes;
    mbar_arrive_expect_tx(&full[sj], planeBytes + (sideL ? G::kSideBytes : 0) + (sideR ? G::kSideBytes : 0));
    tma_load_3d(smem + sj * G::kStageBytes, tmap, cx, cy, z0 + j, &full[sj]);
    /* side boxes: the tile's rows (raw row cy + 1); left tiles want the aligned group of four cells that holds cell
       sx-1 (a TMA box must start on a 16-byte boundary: an unaligned start is an illegal instruction), right tiles
       cells 0 .. 3 */
    if (sideL) tma_load_3d(smem + G::kSideOffset + sj * G::kSideBytes, tmapSide, P.tmaX0 + ((P.sx - 1) & ~3), cy + 1, z0 + j, &full[sj]);
    if (sideR) tma_load_3d(smem + G::kSideOffset + (kTmaStages + sj) * G::kSideBytes, tmapSide, P.tmaX0, cy + 1, z0 + j, &full[sj]);
  };
  if (threadIdx.x == 0)
    for (int j = 0; j < kTmaStages - kTmaLag && j < nplanes; ++j) issue(j);
  {
    /* ---- consumers: a warp owns G::kSub pairs of tile rows, a lane 4 cells of each row of its pair ---- */
    const int r0 = 2 * (warp * G::kSub + sub);
    const int x0 = t.x0 + xl * kVec;
    const int y0 = t.y0 + r0;
    const bool nearSphere = tile_near_sphere(P, t, G::kW, G::kR, z0, z1);
    /* Inside a tile near a sphere only the vectors that really touch it take the (slow, exact) override path. A cell is
       inside when dx^2 + dy^2 + dz^2 < r2; per lane and row, r2 - dy^2 and the smallest dx^2 over the lane's four cells
       and both centres are constants, so each plane costs one multiply and two compares per lane. (The spheres grow with
       the domain: at 2048 cells they cover half of a 2048 x 256 x 256 box, where calling the override for every vector of
       every near tile made the kernel 55 % slower.) */
    int sphA[2] = {0, 0}, sphMinH = 0, sphMaxH = 0, sphMinC = 0, sphMaxC = 0;
    const bool sphFast = nearSphere && P.sphere32;
    if (sphFast) {
      /* squared x distance of the nearest / farthest of the lane's four cells to the hot and the cold centre */
      auto near2 = [](int a0, int a1, int c) { const int d = c < a0 ? a0 - c : (c > a1 ? c - a1 : 0); return d * d; };
      auto far2 = [](int a0, int a1, int c) { const int d = max(abs(a0 - c), abs(a1 - c)); return d * d; };
      const int gx0 = int(P.ox) + x0, gx1 = gx0 + kVec - 1;
      sphMinH = near2(gx0, gx1, int(P.hcx)), sphMaxH = far2(gx0, gx1, int(P.hcx));
      sphMinC = near2(gx0, gx1, int(P.ccx)), sphMaxC = far2(gx0, gx1, int(P.ccx));
      for (int r = 0; r < 2; ++r) {
        const int dy = int(P.oy) + y0 + r - int(P.cy);
        sphA[r] = int(P.r2) - dy * dy;
      }
    }
    /* x wrap by load: which lane's vector holds cell x = sx (patchK = its index there), or is it the last lane's east read? */
    const int patchK = (sideR && x0 <= P.sx && P.sx < x0 + kVec) ? P.sx - x0 : -1;
    const bool eSide = sideR && x0 + kVec == P.sx;
    /* Nothing of the fusion lives in the plane loop: face cells are copied to the neighbours by push_tile_yz() and
       push_tile_x() after it, from what this tile has just written to `next` (still in L2). */
    /* Everything that does not change from plane to plane is decided here, so the plane loop is loads, adds and one
       store per row: where this lane's rows go (running pointers, advanced by one plane per iteration), whether the
       store is a full vector, and whether this lane has anything to push at all. */
    float *dp[2];
    int mode[2]; // 0: nothing to store, 1: full 16-byte store, 2: partial row segment (region edge)
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int y = y0 + r;
      const bool in = y >= P.loy && y < P.hiy && x0 < P.hix && x0 + kVec > P.lox;
      mode[r] = !in ? 0 : ((x0 >= P.lox && x0 + kVec <= P.hix) ? 1 : 2);
      dp[r] = P.dst + (long long)z0 * P.pitchXY + (long long)y * P.pitchX + x0;
    }
    const long long planeStride = P.pitchXY;
    /* Register plan without copies: the centre of plane i is needed again two planes later (as "below"), so two
       register sets alternate (even / odd planes) and the loop is unrolled by two; the in-plane partial sum is
       overwritten in place once the output that needs it has been formed. */
    float4 cmA[2], cmB[2], part[2];
    cmA[0] = cmA[1] = cmB[0] = cmB[1] = part[0] = part[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    int s = 0;
    unsigned phase = 0;
    auto plane = [&](const int i, float4 (&below)[2]) { // `below` holds plane i-2 on entry and plane i on exit
      mbar_wait(&full[s], phase);
      const float *pl = reinterpret_cast<const float *>(smem + s * G::kStageBytes);
      /* smem row j holds compute row t.y0 + j - 1; smem column 4 + k holds compute x = t.x0 + k */
      const float *rowC0 = pl + (r0 + 1) * G::kBoxX + 4 + xl * kVec;
      float4 c0 = *reinterpret_cast<const float4 *>(rowC0);
      float4 c1 = *reinterpret_cast<const float4 *>(rowC0 + G::kBoxX);
      const float *sL = reinterpret_cast<const float *>(smem + G::kSideOffset + s * G::kSideBytes) + r0 * kSideBoxX;
      const float *sR = reinterpret_cast<const float *>(smem + G::kSideOffset + (kTmaStages + s) * G::kSideBytes) + r0 * kSideBoxX;
      if (patchK >= 0) { // my vector contains cell x = sx: it is cell 0 of the row
        const float v0 = sR[0], v1 = sR[kSideBoxX];
        if (patchK == 0) c0.x = v0, c1.x = v1;
        else if (patchK == 1) c0.y = v0, c1.y = v1;
        else if (patchK == 2) c0.z = v0, c1.z = v1;
        else c0.w = v0, c1.w = v1;
      }
      const float4 south = *reinterpret_cast<const float4 *>(rowC0 - G::kBoxX);     // y-1 of row r0
      const float4 north = *reinterpret_cast<const float4 *>(rowC0 + 2 * G::kBoxX); // y+1 of row r1
      float w0 = __shfl_up_sync(0xffffffffu, c0.w, 1, LW), e0 = __shfl_down_sync(0xffffffffu, c0.x, 1, LW);
      float w1 = __shfl_up_sync(0xffffffffu, c1.w, 1, LW), e1 = __shfl_down_sync(0xffffffffu, c1.x, 1, LW);
      if (xl == 0) {
        w0 = sideL ? sL[(P.sx - 1) & 3] : rowC0[-1];
        w1 = sideL ? sL[kSideBoxX + ((P.sx - 1) & 3)] : rowC0[G::kBoxX - 1];
      }
      if (xl == LW - 1) {
        e0 = eSide ? sR[0] : rowC0[kVec];
        e1 = eSide ? sR[kSideBoxX] : rowC0[G::kBoxX + kVec];
      }
      /* this stage's data is in registers now: hand the slot back before doing the arithmetic and the stores */
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      if (threadIdx.x == 0 && i + kTmaStages - kTmaLag < nplanes) issue(i + kTmaStages - kTmaLag);
      if (++s == kTmaStages) {
        s = 0;
        phase ^= 1u;
      }
      if (i >= 2) {
        const int z = z0 + i - 2; // output plane: above = centre of this plane, below = centre two planes back
        float4 o[2];
        o[0].x = div6((part[0].x + c0.x) + below[0].x);
        o[0].y = div6((part[0].y + c0.y) + below[0].y);
        o[0].z = div6((part[0].z + c0.z) + below[0].z);
        o[0].w = div6((part[0].w + c0.w) + below[0].w);
        o[1].x = div6((part[1].x + c1.x) + below[1].x);
        o[1].y = div6((part[1].y + c1.y) + below[1].y);
        o[1].z = div6((part[1].z + c1.z) + below[1].z);
        o[1].w = div6((part[1].w + c1.w) + below[1].w);
        if (sphFast) {
          /* vector wholly inside a sphere: the sphere's temperature; wholly outside both: nothing; else the exact
             per-cell test (cells inside cost 2.7x a normal cell when all of them took that path: a 1024-cell domain
             has 7 % of its cells inside the spheres) */
          const int dz = int(P.oz) + z - int(P.cz), dz2 = dz * dz;
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int rem = sphA[r] - dz2;
            if (sphMinH < rem || sphMinC < rem) {
              if (sphMaxH < rem) {
                const float h = P.dyn[0];
                o[r] = make_float4(h, h, h, h);
              } else if (sphMaxC < rem && !(sphMinH < rem)) {
                const float c = P.dyn[1];
                o[r] = make_float4(c, c, c, c);
              } else {
                o[r] = apply_spheres(P, x0, y0 + r, z, o[r]);
              }
            }
          }
        } else if (nearSphere) {
          o[0] = apply_spheres(P, x0, y0, z, o[0]);
          o[1] = apply_spheres(P, x0, y0 + 1, z, o[1]);
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          if (mode[r] == 1)
            *reinterpret_cast<float4 *>(dp[r]) = o[r];
          else if (mode[r] == 2)
            store_partial(P, dp[r], x0, o[r]);
          dp[r] += planeStride;
        }
      }
      below[0] = c0;
      below[1] = c1;
      /* ((px + mx) + py) + my, the first four terms of the reference's sum, for the plane just read */
      part[0].x = ((c0.y + w0) + c1.x) + south.x;
      part[0].y = ((c0.z + c0.x) + c1.y) + south.y;
      part[0].z = ((c0.w + c0.y) + c1.z) + south.z;
      part[0].w = ((e0 + c0.z) + c1.w) + south.w;
      part[1].x = ((c1.y + w1) + north.x) + c0.x;
      part[1].y = ((c1.z + c1.x) + north.y) + c0.y;
      part[1].z = ((c1.w + c1.y) + north.z) + c0.z;
      part[1].w = ((e1 + c1.z) + north.w) + c0.w;
    };
    /* A warp none of whose lanes has a cell to update (rows past the last row of the subdomain or outside the region:
       645 rows leave 11 of the last block's 16 rows empty) only keeps the ring's protocol going — wait, hand the slot
       back, let thread 0 refill — and leaves the issue slots to the warps that have work. */
    if (__any_sync(0xffffffffu, (mode[0] | mode[1]) != 0)) {
      int i = 0;
      for (; i + 1 < nplanes; i += 2) {
        plane(i, cmA);
        plane(i + 1, cmB);
      }
      if (i < nplanes) plane(i, cmA);
    } else {
      for (int i = 0; i < nplanes; ++i) {
        mbar_wait(&full[s], phase);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (threadIdx.x == 0 && i + kTmaStages - kTmaLag < nplanes) issue(i + kTmaStages - kTmaLag);
        if (++s == kTmaStages) {
          s = 0;
          phase ^= 1u;
        }
      }
    }
  }

  if (FUSED && t.faces) {
    __syncthreads();
    {
      const unsigned yz = (P.debugSkip & 4u) ? 0u : (t.faces & ~((1u << XM) | (1u << XP)));
      if (yz && z1 > z0) push_tile_yz(P, yz, t.x0, t.y0, G::kW, G::kR, z0, z1);
    }
    if (!(P.debugSkip & 2u) && (t.faces & ((1u << XM) | (1u << XP))) && z1 > z0) push_tile_x(P, t.faces, t.y0, G::kR, z0, z1);
    /* only tiles that stored into ANOTHER GPU take part in the completion counts (same-GPU neighbours are ordered by
       the stream, or are this very kernel when an axis wraps onto the subdomain itself) */
    if ((t.faces & P.flagMask) && !(P.debugSkip & 1u)) {
      __syncthreads();
      signal_faces(P, t.faces & P.flagMask);
    }
  }
}

template <bool FUSED, int MINB>
__global__ void __launch_bounds__(kTmaThreads, MINB)
    jacobi7_tma_kernel(const __grid_constant__ CUtensorMap tmapWide, const __grid_constant__ CUtensorMap tmapNarrow,
                       const __grid_constant__ CUtensorMap tmapSideWide, const __grid_constant__ CUtensorMap tmapSideNarrow,
                       const __grid_constant__ JacobiParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); // the next iteration may start filling the slots this grid frees
  const Tile t = P.tiles[blockIdx.x];
  if (t.narrow)
    tma_tile<FUSED, Narrow>(smem, &tmapNarrow, &tmapSideNarrow, P, t);
  else
    tma_tile<FUSED, Wide>(smem, &tmapWide, &tmapSideWide, P, t);
}

__global__ void div6_check_kernel(const float *in, unsigned *mismatches, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float a = div6(in[i]), b = __fdiv_rn(in[i], 6.0f);
  if (__float_as_uint(a) != __float_as_uint(b)) atomicAdd(mismatches, 1u);
}

__global__ void fill_kernel(float *interior, long long pitchX, long long pitchXY, int sx, int sy, int sz, float v) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= sx || y >= sy) return;
  for (int z = blockIdx.z; z < sz; z += gridDim.z) interior[(long long)z * pitchXY + (long long)y * pitchX + x] = v;
}

__host__ __device__ inline unsigned long long mix64(unsigned long long h) {
  h ^= h >> 30;
  h *= 0xBF58476D1CE4E5B9ull;
  h ^= h >> 27;
  h *= 0x94D049BB133111EBull;
  h ^= h >> 31;
  return h;
}
__host__ __device__ inline float random_value(unsigned seed, unsigned long long globalIndex) {
  return float(unsigned(mix64(globalIndex + (unsigned long long)seed * 0x9E3779B97F4A7C15ull) >> 40)) * (1.0f / 16777216.0f);
}

__global__ void spin_kernel(unsigned long long ns) {
  const unsigned long long t0 = globaltimer_ns();
  while (globaltimer_ns() - t0 < ns) __nanosleep(256);
}

__global__ void fill_random_kernel(float *interior, long long pitchX, long long pitchXY, int sx, int sy, int sz, long long ox, long long oy,
                                   long long oz, long long gx, long long gy, unsigned seed) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= sx || y >= sy) return;
  for (int z = blockIdx.z; z < sz; z += gridDim.z)
    interior[(long long)z * pitchXY + (long long)y * pitchX + x] = random_value(seed, (unsigned long long)(((oz + z) * gy + oy + y) * gx + ox + x));
}

__global__ void fingerprint_kernel(const float *interior, long long pitchX, long long pitchXY, int sx, int sy, int sz, long long ox, long long oy,
                                   long long oz, long long gx, long long gy, unsigned long long *out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  unsigned long long acc = 0;
  if (x < sx && y < sy)
    for (int z = blockIdx.z; z < sz; z += gridDim.z) {
      const unsigned long long g = (unsigned long long)(((oz + z) * gy + oy + y) * gx + ox + x);
      const unsigned bits = __float_as_uint(interior[(long long)z * pitchXY + (long long)y * pitchX + x]);
      acc += mix64(((unsigned long long)bits << 32 | 0x5bd1e995u) ^ mix64(g + 1));
    }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && acc) atomicAdd(out, acc);
}

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int d) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != d) CUDA_RUNTIME(cudaSetDevice(d));
  }
  ~DevGuard() { cudaSetDevice(prev); }
};

} // namespace

/* ---------------------------------------------------------------------------------------------------------- */

struct Jacobi3D::Impl {
  struct Dom {
    RcStream stream;
    /* done: end of this subdomain's kernel of the step being enqueued; donePrev: of the previous step (what same-GPU
       neighbours wait for). Two events, swapped every step, so that enqueueing threads never race on one. */
    cudaEvent_t done = nullptr, donePrev = nullptr, t0 = nullptr, t1 = nullptr;
    bool hasDependents = false; // some other local subdomain waits for my `done`
    Tile *tilesAll = nullptr; // whole compute region, tiles that push to another GPU first
    unsigned nAll = 0, nFace[6] = {0, 0, 0, 0, 0, 0};
    Tile *tilesInterior = nullptr;
    unsigned nInterior = 0;
    bool anyNarrow = false; // some tile table of this subdomain holds narrow tiles: launches ask for the larger ring
    std::vector<std::pair<Tile *, unsigned>> tilesExterior;
    std::vector<Rect3> exterior;
    Rect3 interior;
    unsigned *doneCounter = nullptr;
    float *dyn = nullptr; // 2 x kDynBytes of run-time parameters (hot, cold, ...): two slots so that an upload for
                          // the next step never touches what the running kernel reads
    const float *dynNow = nullptr; // the slot the next launch reads
    /* step_io(): uploads and read-backs run on their own streams, chained to the kernels by events, so they overlap
       the neighbouring iterations instead of sitting between two kernels on one stream */
    RcStream ioIn, ioOut;
    cudaEvent_t inReady[2] = {nullptr, nullptr}, kernelDone[2] = {nullptr, nullptr}, outDone[2] = {nullptr, nullptr};
    bool kernelRecorded[2] = {false, false}, outRecorded[2] = {false, false};
    unsigned ioSteps = 0;
    JacobiParams base[2]; // per parity: src = buffer[parity]
    /* per parity: the curr buffer as a 3-D tensor (padded row, rows, planes) with the box of a wide tile, of a narrow
       tile, and the 4-cell-wide boxes at the other end of the rows (x wrap by load) for both shapes */
    CUtensorMap tmap[2], tmapNarrow[2], tmapSide[2], tmapSideNarrow[2];
    bool wrapLoad = false;
    std::vector<int> sameDevDeps;
    WaitSet finalWait;
  };
  std::vector<Dom> doms;
  detail::LaunchPool pool; // one enqueueing thread per local subdomain (1 process x N GPUs)
  unsigned iter = 0;
  unsigned *err = nullptr;
  bool fusedOk = true;
  bool useTma = true;
  int minBlocks = 3; // resident CTAs per SM the TMA kernel is compiled for (when forced through the environment)
  bool minBlocksForced = false;
  int rows = kTmaRows; // tile rows of the selected kernel variant
};

namespace {

/* Tiles covering region [lo,hi) (compute coordinates) of a subdomain of size sz. faceMask selects which subdomain
   faces push (0 for non-fused launches); countMask the faces whose neighbour is another GPU, nFace[f] receives the
   number of tiles that push through face f to such a neighbour. With `narrow` (TMA variant) the part of a row that is
   left after the last full 128-cell tile is covered by 32 x 64 tiles when it is at most kNarrowMaxRemainder cells.

   Order. Natural (z, y, x) order keeps tiles that share halo rows / planes / margins resident together, so those
   re-reads hit L2 (putting ALL pushing tiles first, x faces included, measured +160 MB of DRAM reads per 512^3
   iteration). Tiles that push to another GPU are moved to the front all the same — z faces (whole layers, which keep
   their locality) first, then the y / x face tiles of the other layers in z order: a face is announced to its
   neighbour as soon as its last tile retires, i.e. within the first fifth of the kernel, so the neighbour's next
   kernel — whose first tiles are exactly the ones that need that face — never waits for the END of this one, and the
   GPUs may drift by most of an iteration without anybody spinning. Cost: two planes per z face and one row per y-face
   tile lose their L2 hit (< 1 % of the traffic). */
/* The z layers of planes [lo, hi). The bulk is cut into layers of (about) tileZ planes: deep tiles amortise the two
   halo planes and the pipeline fill of a tile. The top `tailPlanes` planes are cut into layers of tailDepth instead:
   a kernel ends when its slowest CTA does, and with 296 resident tiles of equal depth the last of them retire over
   most of a tile's duration, during which the SMs that finished first have nothing left to take; thin tiles at the
   end of the table fill that gap in small pieces (the same reasoning as for the cheap tiles in make_tiles). Depths
   are even wherever the plane count allows (the plane loop is unrolled by two). */
struct ZPlan {
  int tileZ = 32, tailPlanes = 0, tailDepth = 8;
};
std::vector<std::pair<int64_t, int64_t>> z_layers(int64_t lo, int64_t hi, const ZPlan &zp) {
  std::vector<std::pair<int64_t, int64_t>> out;
  const int64_t planes = hi - lo;
  if (planes <= 0) return out;
  const int64_t tileZ = std::max(1, zp.tileZ), tailDepth = std::max(1, std::min(zp.tailDepth, zp.tileZ));
  /* no thin tail on regions that are not at least two full layers deeper than it */
  int64_t tail = (zp.tailPlanes > 0 && planes >= zp.tailPlanes + 2 * tileZ) ? zp.tailPlanes : 0;
  const int64_t bulk = planes - tail;
  int64_t z = lo;
  if (tail == 0) {
    for (; z < hi; z += tileZ) out.emplace_back(z, std::min(z + tileZ, hi));
    return out;
  }
  /* the bulk in equal (even) layers, so that no short layer ends up in the middle of the table */
  const int64_t n = (bulk + tileZ - 1) / tileZ;
  int64_t d = (bulk + n - 1) / n;
  d += d & 1;
  d = std::min(d, tileZ);
  const int64_t end = lo + bulk;
  for (; z < end; z += d) out.emplace_back(z, std::min(z + d, end));
  for (z = end; z < hi; z += tailDepth) out.emplace_back(z, std::min(z + tailDepth, hi));
  return out;
}

std::vector<Tile> make_tiles(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int rows, const ZPlan &zp, unsigned faceMask, unsigned *nFace,
                             bool narrow, bool flaggedFirst = true, unsigned countMask = 0) {
  const int tileZ = zp.tileZ;
  std::vector<Tile> out;
  if (nFace)
    for (int f = 0; f < 6; ++f) nFace[f] = 0;
  if (hi.x <= lo.x || hi.y <= lo.y || hi.z <= lo.z) return out;
  /* columns of tiles: (x0, narrow?) */
  std::vector<std::pair<int64_t, bool>> cols;
  for (int64_t x = lo.x / kTileX * kTileX; x < hi.x;) {
    const int64_t left = std::min<int64_t>(hi.x, sz.x) - x;
    if (narrow && left <= kNarrowMaxRemainder) {
      /* skip 32-cell columns that lie wholly before lo.x (regions that start inside a tile) */
      if (x + Narrow::kW > lo.x) cols.emplace_back(x, true);
      x += Narrow::kW;
    } else {
      cols.emplace_back(x, false);
      x += kTileX;
    }
  }
  auto emit = [&](const std::pair<int64_t, int64_t> &zl, int64_t x0, int64_t y0, int64_t w, int64_t r, bool nar) {
    unsigned f = 0;
    if (x0 == 0) f |= 1u << XM;
    if (x0 <= sz.x - 1 && sz.x - 1 < x0 + w) f |= 1u << XP;
    if (y0 == 0) f |= 1u << YM;
    if (y0 <= sz.y - 1 && sz.y - 1 < y0 + r) f |= 1u << YP;
    if (zl.first == 0) f |= 1u << ZM;
    if (zl.first <= sz.z - 1 && sz.z - 1 < zl.second) f |= 1u << ZP;
    f &= faceMask;
    if (nFace)
      for (int k = 0; k < 6; ++k)
        if ((f & countMask) >> k & 1u) ++nFace[k];
    out.push_back(Tile{int(x0), int(y0), int(zl.first), (unsigned short)(zl.second - zl.first), (unsigned char)f, (unsigned char)(nar ? 1 : 0)});
  };
  for (const auto &tz : z_layers(lo.z, hi.z, zp)) {
    /* wide columns row block by row block, then the narrow columns (their row blocks are four times as tall) */
    for (int64_t y0 = lo.y / rows * rows; y0 < hi.y; y0 += rows)
      for (const auto &c : cols)
        if (!c.second) emit(tz, c.first, y0, kTileX, rows, false);
    for (int64_t y0 = lo.y / Narrow::kR * Narrow::kR; y0 < hi.y; y0 += Narrow::kR)
      for (const auto &c : cols)
        if (c.second) emit(tz, c.first, y0, Narrow::kW, Narrow::kR, true);
  }
  if (flaggedFirst && countMask) {
    /* ... and the cheap tiles last: a kernel ends with a partial wave of tiles, and what fills it best are tiles that
       take a fraction of a full one — those whose rows are mostly past the last row of the subdomain (their idle warps
       skip the arithmetic) and those of a short top layer. In natural order the short top layer played that part; with
       the z faces moved to the front the tail was all full tiles (2 GPUs, 645 x 645 x 323: 0.2272 against 0.2223 ms). */
    const unsigned zbits = (1u << ZM) | (1u << ZP);
    auto cheap = [&](const Tile &t) {
      const int64_t r = (t.narrow ? Narrow::kR : rows);
      const int64_t activeRows = std::min<int64_t>(t.y0 + r, hi.y) - std::max<int64_t>(t.y0, lo.y);
      return 2 * activeRows <= r || 2 * int64_t(t.nz) <= tileZ;
    };
    auto prio = [&](const Tile &t) { return (t.faces & countMask & zbits) ? 0 : ((t.faces & countMask) ? 1 : (cheap(t) ? 3 : 2)); };
    std::stable_sort(out.begin(), out.end(), [&](const Tile &a, const Tile &b) { return prio(a) < prio(b); });
  }
  return out;
}

} // namespace

namespace detail {
std::vector<TileBox> jacobi_tile_table(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int tileZ, unsigned faceMask, unsigned countMask, bool narrow,
                                       bool flaggedFirst, unsigned nFace[6], int tailPlanes, int tailDepth) {
  std::vector<TileBox> out;
  ZPlan zp;
  zp.tileZ = tileZ;
  zp.tailPlanes = tailPlanes;
  zp.tailDepth = tailDepth;
  for (const Tile &t : make_tiles(lo, hi, sz, kTmaRows, zp, faceMask, nFace, narrow, flaggedFirst, countMask)) {
    const int64_t w = t.narrow ? Narrow::kW : Wide::kW, r = t.narrow ? Narrow::kR : Wide::kR;
    TileBox b;
    b.lo = Dim3(std::max<int64_t>(t.x0, lo.x), std::max<int64_t>(t.y0, lo.y), t.z0);
    b.hi = Dim3(std::min<int64_t>(t.x0 + w, hi.x), std::min<int64_t>(t.y0 + r, hi.y), t.z0 + int64_t(t.nz));
    b.faces = t.faces;
    b.shape = t.narrow;
    b.field = 0;
    out.push_back(b);
  }
  return out;
}
} // namespace detail

namespace {

Tile *upload(const std::vector<Tile> &t) {
  if (t.empty()) return nullptr;
  Tile *d = nullptr;
  CUDA_RUNTIME(cudaMalloc(&d, t.size() * sizeof(Tile)));
  CUDA_RUNTIME(cudaMemcpy(d, t.data(), t.size() * sizeof(Tile), cudaMemcpyHostToDevice));
  return d;
}

/* fp32 tensor over one padded buffer: dims (row pitch, rows, planes), box = boxX x boxY cells of one plane */
CUtensorMap make_plane_map(void *bufferStart, const Dim3 &pitch, const Dim3 &raw, int boxX, int boxY) {
  return tma::make_plane_map(bufferStart, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, sizeof(float), pitch.x, raw.y, raw.z, unsigned(boxX), unsigned(boxY));
}

const Dim3 kFaceDir[6] = {Dim3(-1, 0, 0), Dim3(1, 0, 0), Dim3(0, -1, 0), Dim3(0, 1, 0), Dim3(0, 0, -1), Dim3(0, 0, 1)};

} // namespace

Jacobi3D::Jacobi3D(size_t x, size_t y, size_t z, const Options &opt) : opt_(opt) { build(x, y, z, nullptr); }
Jacobi3D::Jacobi3D(size_t x, size_t y, size_t z, const Options &opt, Comm &comm) : opt_(opt) { build(x, y, z, &comm); }

Jacobi3D::~Jacobi3D() {
  if (impl_) {
    /* a neighbour's last iteration may still be storing into my `next` halo and flags: wait for its final flags and
       let every rank arrive before any IPC-exported slab is freed */
    if (dd_) {
      if (*impl_->err == 0) sync();
      dd_->comm().barrier();
    }
    for (size_t di = 0; di < impl_->doms.size(); ++di) {
      Impl::Dom &d = impl_->doms[di];
      DevGuard g(dd_->domains()[di].gpu());
      cudaStreamSynchronize(d.stream);
      cudaEventDestroy(d.done);
      cudaEventDestroy(d.donePrev);
      cudaEventDestroy(d.t0);
      cudaEventDestroy(d.t1);
      cudaFree(d.tilesAll);
      cudaFree(d.tilesInterior);
      for (auto &p : d.tilesExterior) cudaFree(p.first);
      cudaFree(d.doneCounter);
      cudaFree(d.dyn);
      for (int k = 0; k < 2; ++k) {
        if (d.inReady[k]) cudaEventDestroy(d.inReady[k]);
        if (d.kernelDone[k]) cudaEventDestroy(d.kernelDone[k]);
        if (d.outDone[k]) cudaEventDestroy(d.outDone[k]);
      }
    }
    if (impl_->err) cudaFreeHost(impl_->err);
  }
  impl_.reset();
  dd_.reset();
}

void Jacobi3D::build(size_t x, size_t y, size_t z, Comm *comm) {
  if (opt_.tileZ < 0) opt_.tileZ = 0; // 0: chosen per subdomain below
  if (opt_.useTma && opt_.tileZ > kMaxTileZ) opt_.tileZ = kMaxTileZ;
  if (opt_.tileZ > 32768) opt_.tileZ = 32768; // Tile::nz is 16 bits wide
  if (const char *e = std::getenv("STENCIL_JACOBI_TAIL_PLANES")) opt_.tailPlanes = std::atoi(e);
  if (const char *e = std::getenv("STENCIL_JACOBI_TAIL_DEPTH")) opt_.tailDepth = std::atoi(e);
  if (std::getenv("STENCIL_JACOBI_NATURAL_ORDER")) opt_.flaggedFirst = false;
  if (std::getenv("STENCIL_JACOBI_NO_NARROW")) opt_.narrowTiles = false;
  if (const char *e = std::getenv("STENCIL_JACOBI_PDL")) opt_.pdl = std::atoi(e) != 0;
  if (const char *e = std::getenv("STENCIL_LAUNCH_THREADS")) opt_.launchThreads = std::atoi(e) != 0;
  dd_.reset(comm ? new DistributedDomain(x, y, z, *comm) : new DistributedDomain(x, y, z));
  if (!opt_.gpus.empty()) dd_->set_gpus(opt_.gpus);
  Radius radius = Radius::constant(0);
  radius.set_face(1);
  dd_->set_radius(radius);
  dd_->set_methods(opt_.methods);
  dd_->set_placement(opt_.placement);
  dd_->set_plan_file(opt_.planFile);
  h_ = dd_->add_data<float>("d");
  dd_->realize();

  impl_.reset(new Impl);
  impl_->useTma = opt_.useTma;
  impl_->rows = opt_.useTma ? kTmaRows : kRows;
  if (const char *e = std::getenv("STENCIL_JACOBI_MINB")) {
    impl_->minBlocks = std::atoi(e) == 2 ? 2 : 3;
    impl_->minBlocksForced = true;
  }
  CUDA_RUNTIME(cudaHostAlloc(&impl_->err, sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable));
  *impl_->err = 0;

  const Dim3 global = dd_->size();
  const std::vector<Rect3> interiors = dd_->get_interior();
  const std::vector<std::vector<Rect3>> exteriors = dd_->get_exterior();
  std::vector<LocalDomain> &doms = dd_->domains();
  impl_->doms.resize(doms.size());

  /* can every face neighbour be reached by stores from the local GPU? otherwise fused mode is impossible */
  for (size_t di = 0; di < doms.size() && opt_.mode == Mode::Fused; ++di)
    for (int f = 0; f < 6; ++f) {
      const NeighborInfo n = dd_->neighbor(di, kFaceDir[f], h_.id());
      if (n.exists && !n.reachable) impl_->fusedOk = false;
    }
  {
    /* all ranks must agree on the mode */
    const uint64_t bad = dd_->comm().allreduce_sum(impl_->fusedOk ? 0 : 1);
    if (bad) impl_->fusedOk = false;
  }
  if (opt_.mode == Mode::Fused && !impl_->fusedOk) {
    LOG_WARN("Jacobi3D: a neighbour is not peer-reachable; falling back to the overlapped exchange() schedule");
    opt_.mode = Mode::Overlap;
  }

  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    Impl::Dom &d = impl_->doms[di];
    DevGuard g(ld.gpu());
    d.stream = RcStream(ld.gpu());
    if (opt_.useTma) { // function attributes are per device
      for (auto *k : {&jacobi7_tma_kernel<true, 2>, &jacobi7_tma_kernel<false, 2>, &jacobi7_tma_kernel<true, 3>, &jacobi7_tma_kernel<false, 3>})
        CUDA_RUNTIME(cudaFuncSetAttribute(*k, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes));
      /* three 59 KB rings per SM fit the driver's default carve-out; forcing the maximum carve-out measured no better */
      int carve = -1;
      if (const char *e = std::getenv("STENCIL_JACOBI_CARVEOUT")) carve = std::atoi(e); // percent of the unified array; -1 = driver default
      if (carve >= 0)
        for (auto *k : {&jacobi7_tma_kernel<true, 2>, &jacobi7_tma_kernel<false, 2>, &jacobi7_tma_kernel<true, 3>, &jacobi7_tma_kernel<false, 3>})
          CUDA_RUNTIME(cudaFuncSetAttribute(*k, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    }
    CUDA_RUNTIME(cudaEventCreateWithFlags(&d.done, cudaEventDisableTiming));
    CUDA_RUNTIME(cudaEventCreateWithFlags(&d.donePrev, cudaEventDisableTiming));
    CUDA_RUNTIME(cudaEventCreate(&d.t0));
    CUDA_RUNTIME(cudaEventCreate(&d.t1));
    CUDA_RUNTIME(cudaMalloc(&d.doneCounter, 8 * sizeof(unsigned)));
    CUDA_RUNTIME(cudaMemset(d.doneCounter, 0, 8 * sizeof(unsigned)));
    {
      float init[kDynBytes / sizeof(float)] = {1.0f, 0.0f}; // HOT_TEMP, COLD_TEMP of the reference
      CUDA_RUNTIME(cudaMalloc(&d.dyn, 2 * kDynBytes));
      CUDA_RUNTIME(cudaMemcpy(d.dyn, init, kDynBytes, cudaMemcpyHostToDevice));
      CUDA_RUNTIME(cudaMemcpy(d.dyn + kDynBytes / sizeof(float), init, kDynBytes, cudaMemcpyHostToDevice));
      d.dynNow = d.dyn;
      d.ioIn = RcStream(ld.gpu());
      d.ioOut = RcStream(ld.gpu());
      for (int k = 0; k < 2; ++k) {
        CUDA_RUNTIME(cudaEventCreateWithFlags(&d.inReady[k], cudaEventDisableTiming));
        CUDA_RUNTIME(cudaEventCreateWithFlags(&d.kernelDone[k], cudaEventDisableTiming));
        CUDA_RUNTIME(cudaEventCreateWithFlags(&d.outDone[k], cudaEventDisableTiming));
      }
    }

    const Dim3 sz = ld.size(), org = ld.origin(), pitch = ld.pitch(h_.id());
    unsigned faceMask = 0, flagMask = 0;
    NeighborInfo nbs[6];
    std::set<int> deps;
    d.finalWait.n = 0;
    if (opt_.mode == Mode::Fused)
      for (int f = 0; f < 6; ++f) {
        nbs[f] = dd_->neighbor(di, kFaceDir[f], h_.id());
        if (!nbs[f].exists) continue;
        faceMask |= 1u << f;
        if (!(nbs[f].sameRank && nbs[f].sameDevice)) flagMask |= 1u << f;
        if (nbs[f].sameRank && nbs[f].sameDevice && nbs[f].dom != int(di)) deps.insert(nbs[f].dom);
      }
    /* x wrap by load instead of x halo pushes: both x neighbours are this very subdomain */
    d.wrapLoad = opt_.mode == Mode::Fused && impl_->useTma && !std::getenv("STENCIL_JACOBI_NO_WRAP_LOAD") && nbs[XM].exists &&
                 nbs[XP].exists && nbs[XM].sameRank && nbs[XP].sameRank && nbs[XM].dom == int(di) && nbs[XP].dom == int(di) && sz.x >= kSideBoxX;
    if (d.wrapLoad) faceMask &= ~((1u << XM) | (1u << XP));
    d.sameDevDeps.assign(deps.begin(), deps.end());
    const int rows = impl_->rows;
    const bool narrow = impl_->useTma && opt_.narrowTiles;
    ZPlan zp;
    zp.tileZ = opt_.tileZ;
    if (zp.tileZ < 1) {
      /* automatic depth: deep layers (kAutoTileZ planes) where the subdomain still gives at least four waves of resident
         tiles, shallower ones — down to 16 planes — on small subdomains, which need the tile count more */
      int sms = 148;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ld.gpu());
      const int64_t perLayer = ((sz.x + kTileX - 1) / kTileX) * ((sz.y + rows - 1) / rows);
      int64_t depth = sz.z * perLayer / (4 * 2 * int64_t(sms));
      depth -= depth & 1;
      zp.tileZ = int(std::min<int64_t>(kAutoTileZ, std::max<int64_t>(16, depth)));
    }
    zp.tailPlanes = impl_->useTma ? opt_.tailPlanes : 0;
    zp.tailDepth = opt_.tailDepth;
    const std::vector<Tile> all = make_tiles(Dim3(0, 0, 0), sz, sz, rows, zp, faceMask, d.nFace, narrow, opt_.flaggedFirst, flagMask);
    d.tilesAll = upload(all);
    d.nAll = unsigned(all.size());
    auto note_narrow = [&d](const std::vector<Tile> &v) {
      for (const Tile &t : v) d.anyNarrow = d.anyNarrow || t.narrow;
    };
    note_narrow(all);
    /* every CTA starts by reading its table entry and nothing can be issued before it arrives: keep the table in L2
       across iterations that stream a gigabyte through it (measured 1.4 % on 512^3) */
    if (impl_->useTma) tma::keep_in_l2(d.stream, d.tilesAll, all.size() * sizeof(Tile));
    d.interior = interiors[di];
    d.exterior = exteriors[di];
    const std::vector<Tile> in = make_tiles(d.interior.lo - org, d.interior.hi - org, sz, rows, zp, 0, nullptr, narrow);
    note_narrow(in);
    d.tilesInterior = upload(in);
    d.nInterior = unsigned(in.size());
    for (const Rect3 &r : d.exterior) {
      const std::vector<Tile> ex = make_tiles(r.lo - org, r.hi - org, sz, rows, zp, 0, nullptr, narrow);
      note_narrow(ex);
      d.tilesExterior.emplace_back(upload(ex), unsigned(ex.size()));
    }

    for (int parity = 0; parity < 2; ++parity) {
      JacobiParams &P = d.base[parity];
      std::memset(&P, 0, sizeof(P));
      const long long interiorOff = (pitch.y + 1) * pitch.x + 1; // raw (1,1,1)
      P.src = static_cast<const float *>(ld.buffer(h_.id(), parity)) + interiorOff;
      P.dst = static_cast<float *>(ld.buffer(h_.id(), parity ^ 1)) + interiorOff;
      P.pitchX = pitch.x;
      P.pitchXY = pitch.x * pitch.y;
      P.sx = int(sz.x);
      P.sy = int(sz.y);
      P.sz = int(sz.z);
      P.lox = P.loy = P.loz = 0;
      P.hix = P.sx;
      P.hiy = P.sy;
      P.hiz = P.sz;
      if (const char *e = std::getenv("STENCIL_JACOBI_DEBUG_SKIP")) P.debugSkip = unsigned(std::atoi(e));
      {
        /* logical element (0,0,0) sits `lead` elements into the padded row; compute x=0 is one radius further */
        const char *bufStart = static_cast<const char *>(ld.slab()) + ld.buffer_start_offset(h_.id(), parity);
        P.tmaX0 = int(ld.lead_elems(h_.id())) + 1;
        if (impl_->useTma) {
          char *b = const_cast<char *>(bufStart);
          d.tmap[parity] = make_plane_map(b, pitch, ld.raw_size(), Wide::kBoxX, (P.debugSkip & 8u) ? Wide::kBoxY - 2 : Wide::kBoxY);
          d.tmapNarrow[parity] = make_plane_map(b, pitch, ld.raw_size(), Narrow::kBoxX, Narrow::kBoxY);
          d.tmapSide[parity] = make_plane_map(b, pitch, ld.raw_size(), kSideBoxX, Wide::kR);
          d.tmapSideNarrow[parity] = make_plane_map(b, pitch, ld.raw_size(), kSideBoxX, Narrow::kR);
        }
        P.wrapLoad = d.wrapLoad ? 1 : 0;
      }
      P.ox = org.x;
      P.oy = org.y;
      P.oz = org.z;
      P.hcx = global.x / 3;
      P.ccx = global.x * 2 / 3;
      P.cy = global.y / 2;
      P.cz = global.z / 2;
      const long long R = global.x / 10;
      P.r2 = (R + 1) * (R + 1);
      P.sphere32 = (global.x <= 30000 && global.y <= 30000 && global.z <= 30000) ? 1 : 0;
      P.tiles = d.tilesAll;
      P.dyn = d.dyn;
      for (int f = 0; f < 6; ++f) P.nFace[f] = d.nFace[f];
      P.done = d.doneCounter;
      P.err = impl_->err;
      P.timeoutNs = flag_timeout_ns();
      for (int f = 0; f < 6; ++f) {
        if (!(faceMask & (1u << f))) continue;
        const NeighborInfo &n = nbs[f];
        const BoxView &v = n.view;
        const long long npx = v.pitchY / 4, npxy = v.pitchZ / 4; // elements
        float *nbase = reinterpret_cast<float *>(v.base[parity ^ 1]); // the neighbour's NEXT buffer
        /* address of the neighbour halo cell that receives my compute cell with the face coordinate fixed and the
           two free coordinates = 0; free coordinates then index with the NEIGHBOUR's pitches. raw = compute + 1. */
        long long rx = 1, ry = 1, rz = 1;
        switch (f) {
        case XM: rx = n.size.x + 1; break; // my x=0 -> its +x halo
        case XP: rx = 0; break;            // my x=sx-1 -> its -x halo
        case YM: ry = n.size.y + 1; break;
        case YP: ry = 0; break;
        case ZM: rz = n.size.z + 1; break;
        case ZP: rz = 0; break;
        }
        P.nb[f] = nbase + (rz * npxy + ry * npx + rx);
        P.nbPitchX[f] = npx;
        P.nbPitchXY[f] = npxy;
        const bool viaFlag = !(n.sameRank && n.sameDevice);
        if (viaFlag) {
          P.flagMask |= 1u << f;
          /* I raise the slot "from direction -dir" in the neighbour; the neighbour raises slot "from dir" in me */
          P.nbFlag[f] = n.control + kFusedFlagBase + DirectionMap<int>::slot(kFaceDir[f] * -1);
          P.myFlag[f] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(kFaceDir[f]);
          if (parity == 0) d.finalWait.flag[d.finalWait.n++] = P.myFlag[f];
        }
      }
    }
  }
  for (Impl::Dom &d : impl_->doms)
    for (int dep : d.sameDevDeps) impl_->doms[size_t(dep)].hasDependents = true;
  init();
}

void Jacobi3D::init() { fill(false, 0); }
void Jacobi3D::init_random(unsigned seed) { fill(true, seed); }

void Jacobi3D::fill(bool random, unsigned seed) {
  std::vector<LocalDomain> &doms = dd_->domains();
  /* neighbours may still be storing the last iteration's faces into these buffers */
  if (impl_->iter > 0) {
    sync();
    dd_->comm().barrier();
  }
  const Dim3 global = dd_->size();
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    Impl::Dom &d = impl_->doms[di];
    DevGuard g(ld.gpu());
    const JacobiParams &P = d.base[ld.parity()];
    const dim3 block(128, 4, 1), grid(unsigned((P.sx + 127) / 128), unsigned((P.sy + 3) / 4), unsigned(std::min(P.sz, 64)));
    if (random)
      fill_random_kernel<<<grid, block, 0, d.stream>>>(const_cast<float *>(P.src), P.pitchX, P.pitchXY, P.sx, P.sy, P.sz, P.ox, P.oy, P.oz, global.x,
                                                       global.y, seed);
    else
      fill_kernel<<<grid, block, 0, d.stream>>>(const_cast<float *>(P.src), P.pitchX, P.pitchXY, P.sx, P.sy, P.sz, 0.5f);
    CUDA_RUNTIME(cudaGetLastError());
    ++launches_;
  }
  sync();
  /* halos of curr must be valid before the first fused iteration reads them */
  if (opt_.mode == Mode::Fused) dd_->exchange();
}

void Jacobi3D::step() {
  STENCIL_RANGE("Jacobi3D::step");
  /* a flag wait that timed out inside a kernel only breaks out of its loop: stop here instead of computing further
     iterations on stale halos (the word is host-mapped memory, reading it costs nothing) */
  if (*impl_->err) LOG_FATAL("Jacobi3D: timed out waiting for a neighbour (code " << *impl_->err << ", iteration " << impl_->iter << ")");
  std::vector<LocalDomain> &doms = dd_->domains();
  const size_t nd = doms.size();
  ++impl_->iter;
  auto launch = [&](size_t di, const Tile *tiles, unsigned n, const Rect3 &reg, bool fused) {
    if (!n) return;
    LocalDomain &ld = doms[di];
    Impl::Dom &d = impl_->doms[di];
    JacobiParams P = d.base[ld.parity()];
    const Dim3 org = ld.origin();
    P.tiles = tiles;
    P.lox = int(reg.lo.x - org.x);
    P.loy = int(reg.lo.y - org.y);
    P.loz = int(reg.lo.z - org.z);
    P.hix = int(reg.hi.x - org.x);
    P.hiy = int(reg.hi.y - org.y);
    P.hiz = int(reg.hi.z - org.z);
    P.epoch = impl_->iter;
    P.dyn = d.dynNow;
    if (impl_->useTma) {
      /* the fused variant needs ~100 registers: at 3 CTAs/SM (80 registers) it spills inside the plane loop, and with
         the shared-memory-heavy carve-out those spills miss L1 (measured 4x slower); the plain variant fits in 80 */
      const int mb = impl_->minBlocksForced ? impl_->minBlocks : 2;
      auto *k = fused ? (mb == 3 ? &jacobi7_tma_kernel<true, 3> : &jacobi7_tma_kernel<true, 2>)
                      : (mb == 3 ? &jacobi7_tma_kernel<false, 3> : &jacobi7_tma_kernel<false, 2>);
      const int pa = ld.parity();
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(n, 1, 1);
      cfg.blockDim = dim3(kTmaThreads, 1, 1);
      cfg.dynamicSmemBytes = size_t(d.anyNarrow ? kTmaSmemBytes : Wide::kSmemBytes);
      cfg.stream = d.stream;
      cudaLaunchAttribute attr = {};
      attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr.val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = &attr;
      cfg.numAttrs = opt_.pdl ? 1 : 0;
      CUDA_RUNTIME(cudaLaunchKernelEx(&cfg, *k, d.tmap[pa], d.tmapNarrow[pa], d.tmapSide[pa], d.tmapSideNarrow[pa], P));
    } else if (fused) {
      jacobi7_kernel<true><<<n, kThreads, 0, d.stream>>>(P);
    } else {
      jacobi7_kernel<false><<<n, kThreads, 0, d.stream>>>(P);
    }
    CUDA_RUNTIME(cudaGetLastError());
    ++launches_;
  };

  if (opt_.mode == Mode::Fused) {
    /* same-GPU neighbours (oversubscription) are ordered by events: flags must never be spun on by a kernel that
       shares the device with the kernel that raises them */
    /* each subdomain waits for the PREVIOUS iteration of its same-GPU neighbours (their `done` events were recorded
       in the previous step), launches, and records its own `done`: independent per subdomain, so with several local
       GPUs every subdomain is enqueued from its own host thread */
    const std::function<void(size_t)> enqueue = [&](size_t di) {
      DevGuard g(doms[di].gpu());
      Impl::Dom &d = impl_->doms[di];
      for (int dep : d.sameDevDeps) CUDA_RUNTIME(cudaStreamWaitEvent(d.stream, impl_->doms[size_t(dep)].donePrev, 0));
      launch(di, d.tilesAll, d.nAll, doms[di].get_compute_region(), true);
      if (d.hasDependents) CUDA_RUNTIME(cudaEventRecord(d.done, d.stream));
    };
    if (nd > 1 && opt_.launchThreads)
      impl_->pool.run(nd, enqueue);
    else
      for (size_t di = 0; di < nd; ++di) enqueue(di);
    for (Impl::Dom &d : impl_->doms) std::swap(d.done, d.donePrev);
    dd_->swap();
    return;
  }

  if (opt_.mode == Mode::Overlap)
    for (size_t di = 0; di < nd; ++di) {
      DevGuard g(doms[di].gpu());
      launch(di, impl_->doms[di].tilesInterior, impl_->doms[di].nInterior, impl_->doms[di].interior, false);
    }
  dd_->exchange();
  for (size_t di = 0; di < nd; ++di) {
    DevGuard g(doms[di].gpu());
    Impl::Dom &d = impl_->doms[di];
    if (opt_.mode == Mode::Overlap) {
      for (size_t si = 0; si < d.exterior.size(); ++si) launch(di, d.tilesExterior[si].first, d.tilesExterior[si].second, d.exterior[si], false);
    } else {
      launch(di, d.tilesAll, d.nAll, doms[di].get_compute_region(), false);
    }
  }
  for (size_t di = 0; di < nd; ++di) {
    DevGuard g(doms[di].gpu());
    CUDA_RUNTIME(cudaStreamSynchronize(impl_->doms[di].stream));
  }
  dd_->swap();
}

void Jacobi3D::step_io(const void *paramsPinned, size_t nbytes, float *resultPinned) {
  std::vector<LocalDomain> &doms = dd_->domains();
  if (nbytes > kParamBytes) LOG_FATAL("Jacobi3D::step_io: parameter block larger than " << kParamBytes << " bytes");
  /* inputs of this step go into the parameter slot the previous step is NOT reading, on the upload stream; the
     compute stream waits for the upload. The slot was last read two steps ago. */
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    Impl::Dom &d = impl_->doms[di];
    const int slot = int(d.ioSteps & 1u);
    float *dst = d.dyn + size_t(slot) * (kDynBytes / sizeof(float));
    if (d.kernelRecorded[slot]) CUDA_RUNTIME(cudaStreamWaitEvent(d.ioIn, d.kernelDone[slot], 0));
    CUDA_RUNTIME(cudaMemcpyAsync(dst, paramsPinned, nbytes, cudaMemcpyHostToDevice, d.ioIn));
    CUDA_RUNTIME(cudaEventRecord(d.inReady[slot], d.ioIn));
    CUDA_RUNTIME(cudaStreamWaitEvent(d.stream, d.inReady[slot], 0));
    /* this step's kernel overwrites the buffer whose centre cell the read-back of two steps ago copies */
    if (d.outRecorded[slot]) CUDA_RUNTIME(cudaStreamWaitEvent(d.stream, d.outDone[slot], 0));
    d.dynNow = dst;
  }
  step();
  /* result of this step: the new temperature at the centre of every local compute region, copied on the read-back
     stream once this step's kernel is done (the next step only reads that buffer) */
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    DevGuard g(ld.gpu());
    Impl::Dom &d = impl_->doms[di];
    const int slot = int(d.ioSteps & 1u);
    CUDA_RUNTIME(cudaEventRecord(d.kernelDone[slot], d.stream));
    d.kernelRecorded[slot] = true;
    CUDA_RUNTIME(cudaStreamWaitEvent(d.ioOut, d.kernelDone[slot], 0));
    const Rect3 r = ld.get_compute_region();
    const Dim3 centre((r.lo.x + r.hi.x) / 2, (r.lo.y + r.hi.y) / 2, (r.lo.z + r.hi.z) / 2);
    Accessor<float> acc = ld.get_curr_accessor(h_);
    CUDA_RUNTIME(cudaMemcpyAsync(resultPinned + di, &acc[centre], sizeof(float), cudaMemcpyDeviceToHost, d.ioOut));
    CUDA_RUNTIME(cudaEventRecord(d.outDone[slot], d.ioOut));
    d.outRecorded[slot] = true;
    ++d.ioSteps;
  }
}

void Jacobi3D::sync() {
  std::vector<LocalDomain> &doms = dd_->domains();
  const unsigned long long timeoutNs = flag_timeout_ns();
  const std::function<void(size_t)> drain = [&](size_t di) {
    DevGuard g(doms[di].gpu());
    Impl::Dom &d = impl_->doms[di];
    /* fused: my halos of the final iteration are complete only when every neighbour raised its last flag */
    if (opt_.mode == Mode::Fused && impl_->iter > 0) launch_wait_flags(d.finalWait, impl_->iter, impl_->err, timeoutNs, d.stream);
    CUDA_RUNTIME(cudaStreamSynchronize(d.stream));
    CUDA_RUNTIME(cudaStreamSynchronize(d.ioOut));
  };
  /* several local GPUs: wait for all of them at once, not one after the other */
  if (doms.size() > 1 && opt_.launchThreads)
    impl_->pool.run(doms.size(), drain);
  else
    for (size_t di = 0; di < doms.size(); ++di) drain(di);
  if (*impl_->err) LOG_FATAL("Jacobi3D: timed out waiting for a neighbour (code " << *impl_->err << ", iteration " << impl_->iter << ")");
}

double Jacobi3D::run(int iters) {
  std::vector<LocalDomain> &doms = dd_->domains();
  sync();
  dd_->comm().barrier();
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    CUDA_RUNTIME(cudaEventRecord(impl_->doms[di].t0, impl_->doms[di].stream));
  }
  for (int i = 0; i < iters; ++i) step();
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    CUDA_RUNTIME(cudaEventRecord(impl_->doms[di].t1, impl_->doms[di].stream));
  }
  sync();
  float worst = 0;
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    float ms = 0;
    CUDA_RUNTIME(cudaEventElapsedTime(&ms, impl_->doms[di].t0, impl_->doms[di].t1));
    worst = std::max(worst, ms);
  }
  return dd_->comm().allreduce_max(double(worst) * 1e-3);
}

double Jacobi3D::checksum() {
  sync();
  std::vector<LocalDomain> &doms = dd_->domains();
  double total = 0;
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    DevGuard g(ld.gpu());
    const JacobiParams &P = impl_->doms[di].base[ld.parity()];
    const double h = reduce_sum_f32_sync(P.src, Dim3(P.pitchX, P.pitchXY / P.pitchX, 1), Dim3(P.sx, P.sy, P.sz), impl_->doms[di].stream);
    total += h;
  }
  /* sum over ranks in a fixed order */
  std::vector<double> all(static_cast<size_t>(dd_->comm().size()));
  dd_->comm().allgather(&total, all.data(), sizeof(double));
  double s = 0;
  for (double v : all) s += v;
  return s;
}

void Jacobi3D::delay(size_t di, unsigned long long ns) {
  if (di >= impl_->doms.size()) return;
  DevGuard g(dd_->domains()[di].gpu());
  spin_kernel<<<1, 1, 0, impl_->doms[di].stream>>>(ns);
  CUDA_RUNTIME(cudaGetLastError());
}

unsigned long long Jacobi3D::fingerprint() {
  sync();
  std::vector<LocalDomain> &doms = dd_->domains();
  const Dim3 global = dd_->size();
  unsigned long long total = 0;
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    Impl::Dom &d = impl_->doms[di];
    DevGuard g(ld.gpu());
    const JacobiParams &P = d.base[ld.parity()];
    /* the per-face counters are idle between iterations; word 6 of that block is free: use words 6..7 as the accumulator */
    unsigned long long *acc = reinterpret_cast<unsigned long long *>(d.doneCounter + 6);
    CUDA_RUNTIME(cudaMemsetAsync(acc, 0, sizeof(unsigned long long), d.stream));
    const dim3 block(128, 4, 1), grid(unsigned((P.sx + 127) / 128), unsigned((P.sy + 3) / 4), unsigned(std::min(P.sz, 64)));
    fingerprint_kernel<<<grid, block, 0, d.stream>>>(P.src, P.pitchX, P.pitchXY, P.sx, P.sy, P.sz, P.ox, P.oy, P.oz, global.x, global.y, acc);
    CUDA_RUNTIME(cudaGetLastError());
    unsigned long long h = 0;
    CUDA_RUNTIME(cudaMemcpyAsync(&h, acc, sizeof(h), cudaMemcpyDeviceToHost, d.stream));
    CUDA_RUNTIME(cudaStreamSynchronize(d.stream));
    total += h;
    ++launches_;
  }
  std::vector<unsigned long long> all(static_cast<size_t>(dd_->comm().size()));
  dd_->comm().allgather(&total, all.data(), sizeof(total));
  unsigned long long s = 0;
  for (unsigned long long v : all) s += v;
  return s;
}

float jacobi_random_value(unsigned seed, const Dim3 &p, const Dim3 &global) {
  return random_value(seed, (unsigned long long)((p.z * global.y + p.y) * global.x + p.x));
}

bool Jacobi3D::probe(const Dim3 &p, float *out) {
  sync();
  for (LocalDomain &ld : dd_->domains()) {
    if (!ld.get_compute_region().contains(p)) continue;
    DevGuard g(ld.gpu());
    Accessor<float> acc = ld.get_curr_accessor(h_);
    CUDA_RUNTIME(cudaMemcpy(out, &acc[p], sizeof(float), cudaMemcpyDeviceToHost));
    return true;
  }
  return false;
}

unsigned jacobi_div6_mismatches(const std::vector<float> &values) {
  float *d = nullptr;
  unsigned *m = nullptr, h = 0;
  CUDA_RUNTIME(cudaMalloc(&d, values.size() * sizeof(float)));
  CUDA_RUNTIME(cudaMalloc(&m, sizeof(unsigned)));
  CUDA_RUNTIME(cudaMemset(m, 0, sizeof(unsigned)));
  CUDA_RUNTIME(cudaMemcpy(d, values.data(), values.size() * sizeof(float), cudaMemcpyHostToDevice));
  div6_check_kernel<<<unsigned((values.size() + 255) / 256), 256>>>(d, m, int(values.size()));
  CUDA_RUNTIME(cudaMemcpy(&h, m, sizeof(unsigned), cudaMemcpyDeviceToHost));
  CUDA_RUNTIME(cudaFree(d));
  CUDA_RUNTIME(cudaFree(m));
  return h;
}

void jacobi_host_step(std::vector<float> &next, const std::vector<float> &curr, const Dim3 &size) {
  const int64_t X = size.x, Y = size.y, Z = size.z;
  const int64_t hcx = X / 3, ccx = X * 2 / 3, cy = Y / 2, cz = Z / 2, R = X / 10, r2 = (R + 1) * (R + 1);
  next.resize(curr.size());
  auto at = [&](int64_t x, int64_t y, int64_t z) { return curr[size_t((((z + Z) % Z) * Y + (y + Y) % Y) * X + (x + X) % X)]; };
  for (int64_t z = 0; z < Z; ++z)
    for (int64_t y = 0; y < Y; ++y)
      for (int64_t x = 0; x < X; ++x) {
        const int64_t b = (y - cy) * (y - cy) + (z - cz) * (z - cz);
        float v;
        if (b + (x - hcx) * (x - hcx) < r2) {
          v = 1.0f;
        } else if (b + (x - ccx) * (x - ccx) < r2) {
          v = 0.0f;
        } else {
          v = 0;
          v += at(x + 1, y, z);
          v += at(x - 1, y, z);
          v += at(x, y + 1, z);
          v += at(x, y - 1, z);
          v += at(x, y, z + 1);
          v += at(x, y, z - 1);
          v /= 6;
        }
        next[size_t((z * Y + y) * X + x)] = v;
      }
}

} // namespace stencil
