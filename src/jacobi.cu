/* Jacobi3D engine and its sm_100a kernels (N2 jacobi7 fused exterior, N3 jacobi7 interior, N9 init).
   Replaces reference kernels K8/K9 (bin/jacobi3d.cu:18-87). See include/stencil/jacobi.hpp for the modes.

   Kernel shape. A CTA is 32 lanes x 8 rows; a lane owns 4 consecutive x cells (one 16-byte vector), so a tile is
   128 x 8 cells in xy and is marched along z for tileZ planes with the centre column kept in registers
   (below / cur / above rotate). Per plane a thread issues three 16-byte loads (above, +y row, -y row); the x
   neighbours come from the adjacent lanes by shuffle, only lane 0 / lane 31 load one scalar each. Rows start on
   128-byte boundaries (LocalDomain layout), so every vector access is aligned and a warp reads whole cache lines.
   HBM traffic is one read + one write of the field per iteration; the +-y / +-z re-reads are served by L1/L2.

   Fusion. Tiles that own cells on a subdomain face additionally store those cells into the neighbour's `next`
   halo (peer pointer: NVLink store, or a plain local store when the axis wraps onto the same subdomain) —
   16-byte vectors for y and z faces, one 4-byte store per row for x faces. Boundary tiles come first in the tile
   table; each finishes with fence.sys + a counter increment, and the last one publishes the iteration number to
   every neighbour's flag (st.release.sys). At the start of the next iteration only boundary tiles look at the
   flags (ld.acquire.sys): they are the only ones that read halo cells, and their neighbours' flags were raised
   well before the previous kernel ended, so in steady state nobody spins. */
#include "stencil/jacobi.hpp"

#include <algorithm>
#include <cstring>
#include <set>

#include "stencil/copy_plan.hpp"
#include "stencil/reduce.hpp"

namespace stencil {

namespace {

constexpr int kLanes = 32, kRows = 8, kThreads = kLanes * kRows, kVec = 4, kTileX = kLanes * kVec;
constexpr size_t kDynBytes = Jacobi3D::kParamBytes;
constexpr int kFusedFlagBase = 64; // word offset of the fused-iteration flags inside a control block

enum Face { XM = 0, XP = 1, YM = 2, YP = 3, ZM = 4, ZP = 5 };

struct Tile {
  unsigned short tx, ty, tz, faces; // faces: bit f set when the tile owns cells on face f that must be pushed
};

struct JacobiParams {
  const float *src; // interior origin: compute cell (0,0,0) of curr
  float *dst;       // interior origin of next
  long long pitchX, pitchXY; // elements: row pitch, plane pitch
  int sx, sy, sz;            // compute-region size
  int lox, loy, loz, hix, hiy, hiz; // cells to update (compute coordinates, half open)
  int tileZ;
  /* Dirichlet spheres in global coordinates */
  long long ox, oy, oz;   // global coordinate of compute cell (0,0,0)
  long long hcx, ccx, cy, cz, r2; // hot/cold centre x, shared centre y/z, (R+1)^2
  const Tile *tiles;
  const float *dyn; // device-resident run-time parameters: [0] hot temperature, [1] cold temperature
  /* fused push */
  float *nb[6];            // neighbour's next buffer: address of ITS halo cell that receives my face cell (0,0) ...
  long long nbPitchX[6], nbPitchXY[6];
  unsigned *nbFlag[6];     // neighbour flag to raise (nullptr: ordered by the stream / same kernel)
  const unsigned *myFlag[6]; // my flag the neighbour raises (nullptr: no wait)
  unsigned epoch;          // iteration number being computed (1-based)
  unsigned nBoundary;      // number of tiles that push
  unsigned *done;          // boundary tile counter
  unsigned *err;           // host-mapped error word
  unsigned long long timeoutNs;
};

/* Plain (coherent) loads on purpose: halo cells are written by OTHER GPUs while this kernel may already be running,
   and only become readable after the acquire on the neighbour's flag. ld.global.nc is outside that guarantee. */
__device__ __forceinline__ float4 ldg4(const float *p) { return *reinterpret_cast<const float4 *>(p); }
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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

/* one output value, in the reference's summation order, with the sphere override */
__device__ __forceinline__ float relax(float px, float mx, float py, float my, float pz, float mz) {
  float v = 0;
  v += px;
  v += mx;
  v += py;
  v += my;
  v += pz;
  v += mz;
  return __fdiv_rn(v, 6.0f);
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

template <bool FUSED> __global__ void __launch_bounds__(kThreads) jacobi7_kernel(const __grid_constant__ JacobiParams P) {
  const Tile t = P.tiles[blockIdx.x];
  const int lane = threadIdx.x & (kLanes - 1), row = threadIdx.x / kLanes;

  if (FUSED && t.faces) {
    /* boundary tiles read halo cells: wait until every neighbour has delivered iteration epoch-1 */
    if (threadIdx.x < 6) {
      const unsigned *f = P.myFlag[threadIdx.x];
      if (f) {
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
    __syncthreads();
  }

  const int x0 = t.tx * kTileX + lane * kVec;
  const int y = t.ty * kRows + row;
  const int z0 = max(int(t.tz) * P.tileZ, P.loz), z1 = min(int(t.tz + 1) * P.tileZ, P.hiz);
  /* a warp is one row, so the row test is warp-uniform; the x test is per lane (partial tiles) */
  const bool rowActive = y >= P.loy && y < P.hiy;
  const bool laneActive = x0 < P.hix && x0 + kVec > P.lox;

  /* does this tile come near either sphere? (uniform per CTA) */
  bool nearSphere;
  {
    const long long gy0 = P.oy + t.ty * kRows, gy1 = gy0 + kRows - 1;
    const long long gz0 = P.oz + z0, gz1 = P.oz + z1 - 1;
    const long long gx0 = P.ox + t.tx * kTileX, gx1 = gx0 + kTileX - 1;
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

  if (FUSED && t.faces) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      const unsigned prev = atomicAdd(P.done, 1u);
      if (prev == P.nBoundary - 1) {
        *P.done = 0;
        __threadfence_system();
        for (int f = 0; f < 6; ++f)
          if (P.nbFlag[f]) st_release_sys(P.nbFlag[f], P.epoch);
      }
    }
  }
}

__global__ void fill_kernel(float *interior, long long pitchX, long long pitchXY, int sx, int sy, int sz, float v) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= sx || y >= sy) return;
  for (int z = blockIdx.z; z < sz; z += gridDim.z) interior[(long long)z * pitchXY + (long long)y * pitchX + x] = v;
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
    cudaEvent_t done = nullptr, t0 = nullptr, t1 = nullptr;
    Tile *tilesAll = nullptr; // whole compute region, boundary tiles first
    unsigned nAll = 0, nBoundary = 0;
    Tile *tilesInterior = nullptr;
    unsigned nInterior = 0;
    std::vector<std::pair<Tile *, unsigned>> tilesExterior;
    std::vector<Rect3> exterior;
    Rect3 interior;
    unsigned *doneCounter = nullptr;
    float *dyn = nullptr; // kDynBytes of run-time parameters (hot, cold, ...)
    JacobiParams base[2]; // per parity: src = buffer[parity]
    std::vector<int> sameDevDeps;
    WaitSet finalWait;
  };
  std::vector<Dom> doms;
  unsigned iter = 0;
  unsigned *err = nullptr;
  bool fusedOk = true;
};

namespace {

/* tiles covering region [lo,hi) (compute coordinates) of a subdomain of size sz; faceMask selects which subdomain
   faces push (0 for non-fused launches). Boundary tiles are emitted first. */
std::vector<Tile> make_tiles(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int tileZ, unsigned faceMask, unsigned *nBoundary) {
  std::vector<Tile> boundary, inner;
  if (hi.x <= lo.x || hi.y <= lo.y || hi.z <= lo.z) {
    if (nBoundary) *nBoundary = 0;
    return inner;
  }
  for (int64_t tz = lo.z / tileZ; tz * tileZ < hi.z; ++tz)
    for (int64_t ty = lo.y / kRows; ty * kRows < hi.y; ++ty)
      for (int64_t tx = lo.x / kTileX; tx * kTileX < hi.x; ++tx) {
        unsigned f = 0;
        if (tx == 0) f |= 1u << XM;
        if (tx * kTileX <= sz.x - 1 && sz.x - 1 < (tx + 1) * kTileX) f |= 1u << XP;
        if (ty == 0) f |= 1u << YM;
        if (ty * kRows <= sz.y - 1 && sz.y - 1 < (ty + 1) * kRows) f |= 1u << YP;
        if (tz == 0) f |= 1u << ZM;
        if (tz * tileZ <= sz.z - 1 && sz.z - 1 < (tz + 1) * tileZ) f |= 1u << ZP;
        f &= faceMask;
        const Tile t{(unsigned short)tx, (unsigned short)ty, (unsigned short)tz, (unsigned short)f};
        (f ? boundary : inner).push_back(t);
      }
  if (nBoundary) *nBoundary = unsigned(boundary.size());
  boundary.insert(boundary.end(), inner.begin(), inner.end());
  return boundary;
}

Tile *upload(const std::vector<Tile> &t) {
  if (t.empty()) return nullptr;
  Tile *d = nullptr;
  CUDA_RUNTIME(cudaMalloc(&d, t.size() * sizeof(Tile)));
  CUDA_RUNTIME(cudaMemcpy(d, t.data(), t.size() * sizeof(Tile), cudaMemcpyHostToDevice));
  return d;
}

const Dim3 kFaceDir[6] = {Dim3(-1, 0, 0), Dim3(1, 0, 0), Dim3(0, -1, 0), Dim3(0, 1, 0), Dim3(0, 0, -1), Dim3(0, 0, 1)};

} // namespace

Jacobi3D::Jacobi3D(size_t x, size_t y, size_t z, const Options &opt) : opt_(opt) { build(x, y, z, nullptr); }
Jacobi3D::Jacobi3D(size_t x, size_t y, size_t z, const Options &opt, Comm &comm) : opt_(opt) { build(x, y, z, &comm); }

Jacobi3D::~Jacobi3D() {
  if (impl_) {
    for (size_t di = 0; di < impl_->doms.size(); ++di) {
      Impl::Dom &d = impl_->doms[di];
      DevGuard g(dd_->domains()[di].gpu());
      cudaStreamSynchronize(d.stream);
      cudaEventDestroy(d.done);
      cudaEventDestroy(d.t0);
      cudaEventDestroy(d.t1);
      cudaFree(d.tilesAll);
      cudaFree(d.tilesInterior);
      for (auto &p : d.tilesExterior) cudaFree(p.first);
      cudaFree(d.doneCounter);
      cudaFree(d.dyn);
    }
    if (impl_->err) cudaFreeHost(impl_->err);
  }
  impl_.reset();
  dd_.reset();
}

void Jacobi3D::build(size_t x, size_t y, size_t z, Comm *comm) {
  if (opt_.tileZ < 1) opt_.tileZ = 32;
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
    CUDA_RUNTIME(cudaEventCreateWithFlags(&d.done, cudaEventDisableTiming));
    CUDA_RUNTIME(cudaEventCreate(&d.t0));
    CUDA_RUNTIME(cudaEventCreate(&d.t1));
    CUDA_RUNTIME(cudaMalloc(&d.doneCounter, sizeof(unsigned)));
    CUDA_RUNTIME(cudaMemset(d.doneCounter, 0, sizeof(unsigned)));
    {
      float init[kDynBytes / sizeof(float)] = {1.0f, 0.0f}; // HOT_TEMP, COLD_TEMP of the reference
      CUDA_RUNTIME(cudaMalloc(&d.dyn, kDynBytes));
      CUDA_RUNTIME(cudaMemcpy(d.dyn, init, kDynBytes, cudaMemcpyHostToDevice));
    }

    const Dim3 sz = ld.size(), org = ld.origin(), pitch = ld.pitch(h_.id());
    unsigned faceMask = 0;
    NeighborInfo nbs[6];
    std::set<int> deps;
    d.finalWait.n = 0;
    if (opt_.mode == Mode::Fused)
      for (int f = 0; f < 6; ++f) {
        nbs[f] = dd_->neighbor(di, kFaceDir[f], h_.id());
        if (!nbs[f].exists) continue;
        faceMask |= 1u << f;
        if (nbs[f].sameRank && nbs[f].sameDevice && nbs[f].dom != int(di)) deps.insert(nbs[f].dom);
      }
    d.sameDevDeps.assign(deps.begin(), deps.end());

    const std::vector<Tile> all = make_tiles(Dim3(0, 0, 0), sz, sz, opt_.tileZ, faceMask, &d.nBoundary);
    d.tilesAll = upload(all);
    d.nAll = unsigned(all.size());
    d.interior = interiors[di];
    d.exterior = exteriors[di];
    const std::vector<Tile> in = make_tiles(d.interior.lo - org, d.interior.hi - org, sz, opt_.tileZ, 0, nullptr);
    d.tilesInterior = upload(in);
    d.nInterior = unsigned(in.size());
    for (const Rect3 &r : d.exterior) {
      const std::vector<Tile> ex = make_tiles(r.lo - org, r.hi - org, sz, opt_.tileZ, 0, nullptr);
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
      P.tileZ = opt_.tileZ;
      P.ox = org.x;
      P.oy = org.y;
      P.oz = org.z;
      P.hcx = global.x / 3;
      P.ccx = global.x * 2 / 3;
      P.cy = global.y / 2;
      P.cz = global.z / 2;
      const long long R = global.x / 10;
      P.r2 = (R + 1) * (R + 1);
      P.tiles = d.tilesAll;
      P.dyn = d.dyn;
      P.nBoundary = d.nBoundary;
      P.done = d.doneCounter;
      P.err = impl_->err;
      P.timeoutNs = 5000000000ull;
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
          /* I raise the slot "from direction -dir" in the neighbour; the neighbour raises slot "from dir" in me */
          P.nbFlag[f] = n.control + kFusedFlagBase + DirectionMap<int>::slot(kFaceDir[f] * -1);
          P.myFlag[f] = ld.arrive_flags() + kFusedFlagBase + DirectionMap<int>::slot(kFaceDir[f]);
          if (parity == 0) d.finalWait.flag[d.finalWait.n++] = P.myFlag[f];
        }
      }
    }
  }
  init();
}

void Jacobi3D::init() {
  std::vector<LocalDomain> &doms = dd_->domains();
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    Impl::Dom &d = impl_->doms[di];
    DevGuard g(ld.gpu());
    const JacobiParams &P = d.base[ld.parity()];
    const dim3 block(128, 4, 1), grid(unsigned((P.sx + 127) / 128), unsigned((P.sy + 3) / 4), unsigned(std::min(P.sz, 64)));
    fill_kernel<<<grid, block, 0, d.stream>>>(const_cast<float *>(P.src), P.pitchX, P.pitchXY, P.sx, P.sy, P.sz, 0.5f);
    CUDA_RUNTIME(cudaGetLastError());
    ++launches_;
  }
  sync();
  /* halos of curr must be valid before the first fused iteration reads them */
  if (opt_.mode == Mode::Fused) dd_->exchange();
}

void Jacobi3D::step() {
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
    if (fused)
      jacobi7_kernel<true><<<n, kThreads, 0, d.stream>>>(P);
    else
      jacobi7_kernel<false><<<n, kThreads, 0, d.stream>>>(P);
    CUDA_RUNTIME(cudaGetLastError());
    ++launches_;
  };

  if (opt_.mode == Mode::Fused) {
    /* same-GPU neighbours (oversubscription) are ordered by events: flags must never be spun on by a kernel that
       shares the device with the kernel that raises them */
    for (size_t di = 0; di < nd; ++di) {
      DevGuard g(doms[di].gpu());
      for (int dep : impl_->doms[di].sameDevDeps) CUDA_RUNTIME(cudaStreamWaitEvent(impl_->doms[di].stream, impl_->doms[size_t(dep)].done, 0));
    }
    for (size_t di = 0; di < nd; ++di) {
      DevGuard g(doms[di].gpu());
      launch(di, impl_->doms[di].tilesAll, impl_->doms[di].nAll, doms[di].get_compute_region(), true);
      if (!impl_->doms[di].sameDevDeps.empty() || nd > 1) CUDA_RUNTIME(cudaEventRecord(impl_->doms[di].done, impl_->doms[di].stream));
    }
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
  /* inputs of this step: stream-ordered before the kernels that read them */
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    CUDA_RUNTIME(cudaMemcpyAsync(impl_->doms[di].dyn, paramsPinned, nbytes, cudaMemcpyHostToDevice, impl_->doms[di].stream));
  }
  step();
  /* result of this step: the new temperature at the centre of every local compute region */
  for (size_t di = 0; di < doms.size(); ++di) {
    LocalDomain &ld = doms[di];
    DevGuard g(ld.gpu());
    const Rect3 r = ld.get_compute_region();
    const Dim3 centre((r.lo.x + r.hi.x) / 2, (r.lo.y + r.hi.y) / 2, (r.lo.z + r.hi.z) / 2);
    Accessor<float> acc = ld.get_curr_accessor(h_);
    CUDA_RUNTIME(cudaMemcpyAsync(resultPinned + di, &acc[centre], sizeof(float), cudaMemcpyDeviceToHost, impl_->doms[di].stream));
  }
}

void Jacobi3D::sync() {
  std::vector<LocalDomain> &doms = dd_->domains();
  for (size_t di = 0; di < doms.size(); ++di) {
    DevGuard g(doms[di].gpu());
    Impl::Dom &d = impl_->doms[di];
    /* fused: my halos of the final iteration are complete only when every neighbour raised its last flag */
    if (opt_.mode == Mode::Fused && impl_->iter > 0) launch_wait_flags(d.finalWait, impl_->iter, impl_->err, 5000000000ull, d.stream);
    CUDA_RUNTIME(cudaStreamSynchronize(d.stream));
  }
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
