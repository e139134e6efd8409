#pragma once
/* Jacobi3D: the 7-point Jacobi heat solver of the reference's bin/jacobi3d.cu (fp32, face radius 1, hot and cold
   Dirichlet spheres standing in for boundary conditions, periodic domain) as a reusable engine, so the CLI app,
   the C ABI / Python binding and bench.py all run the same code.

   Reference behaviour kept (bin/jacobi3d.cu:18-87, :205-352): init to (HOT+COLD)/2 = 0.5; a cell within
   sphereRadius = X/10 of the hot centre (X/3, Y/2, Z/2) is 1, of the cold centre (2X/3, Y/2, Z/2) is 0, where the
   distance is truncated to an integer; otherwise the cell becomes (+x + -x + +y + -y + +z + -z) / 6 in exactly that
   summation order (bit-exact against a host loop). Three execution modes:

     Mode::Fused     (default) ONE kernel per subdomain per iteration computes the whole compute region; tiles on a
                     subdomain face also store their results straight into the neighbour GPU's `next` halo over
                     NVLink and the last boundary tile raises the neighbour's epoch flag. No exchange() call, no
                     host synchronisation inside the loop; the stores to the neighbours are spread over the whole
                     kernel, so the transfer hides behind the computation. (N2 + N3.)
     Mode::Overlap   the reference's schedule: interior kernel, exchange(), exterior slabs, synchronise, swap.
     Mode::NoOverlap exchange(), whole-region kernel, synchronise, swap (reference --no-overlap). */
#include <atomic>
#include <memory>
#include <vector>

#include "stencil/stencil.hpp"

namespace stencil {

class Jacobi3D {
public:
  enum class Mode { Fused, Overlap, NoOverlap };
  struct Options {
    MethodFlags methods = MethodFlags::All;
    PlacementStrategy placement = PlacementStrategy::NodeAware;
    Mode mode = Mode::Fused;
    std::vector<int> gpus; // empty: the rank's default devices
    int tileZ = 0;         // z-extent of a tile (planes marched per CTA); 0: 48 on large subdomains, down to 16 on small ones
    int tailPlanes = 48;   // the top tailPlanes planes of a subdomain are cut into layers of tailDepth planes instead: thin
    int tailDepth = 8;     // tiles at the end of the table fill the kernel's ragged last wave (STENCIL_JACOBI_TAIL_PLANES / _DEPTH)
    bool useTma = true;    // TMA + mbarrier pipelined kernel (false: plain LDG variant, kept for comparison)
    bool flaggedFirst = true; // schedule the tiles that push to ANOTHER GPU first, so each face is announced early
    bool narrowTiles = true;  // cover the remainder of rows that are not a multiple of 128 cells with 32 x 64 tiles
    bool pdl = true;          // programmatic dependent launch between consecutive kernels of a stream: the next iteration's
                              // CTAs take the slots the last wave frees and set up while it drains (STENCIL_JACOBI_PDL=0: off)
    bool planFile = false;
    bool launchThreads = true; // several local GPUs: enqueue each GPU's work from its own host thread (STENCIL_LAUNCH_THREADS=0 turns it off)
  };

  Jacobi3D(size_t x, size_t y, size_t z, const Options &opt);
  Jacobi3D(size_t x, size_t y, size_t z, const Options &opt, Comm &comm);
  ~Jacobi3D();

  DistributedDomain &domain() { return *dd_; }
  const DataHandle<float> &handle() const { return h_; }
  Mode mode() const { return opt_.mode; }

  /* curr = 0.5 on every compute region (halos are filled by the first exchange / fused step); also resets the
     iteration count of the engine's flags hand-shake consistently on all ranks (collective) */
  void init();
  /* curr = a pseudo-random value in [0,1) that depends only on `seed` and the cell's GLOBAL coordinate
     (jacobi_random_value), so every partition and every schedule starts from the same field (collective) */
  void init_random(unsigned seed);
  /* enqueue one iteration and swap. Fused mode does not block the host. */
  void step();
  /* One iteration with per-step host I/O, all stream-ordered (no host blocking): first copy `nbytes`
     (<= kParamBytes) of run-time parameters from PINNED host memory to every local GPU — floats [0] hot sphere
     temperature, [1] cold sphere temperature — then step(), then copy the new value at the centre of each local
     compute region to resultPinned[di] (pinned). Call sync() before reading the results. */
  static constexpr size_t kParamBytes = 64;
  void step_io(const void *paramsPinned, size_t nbytes, float *resultPinned);
  /* block until every local GPU finished all enqueued iterations */
  void sync();
  /* run `iters` iterations and return the device time in seconds: CUDA events per local GPU around the whole
     batch, max over GPUs and over ranks */
  double run(int iters);
  /* kernels enqueued by this object so far (all local GPUs) */
  unsigned long long launches() const { return launches_; }
  /* test hook: keep the stream of local subdomain di busy for about `ns` nanoseconds before whatever is enqueued
     next, so that this GPU falls behind its neighbours (opens the window in which a fast neighbour's pushes of
     iteration k+1 could overtake this GPU's reads of iteration k if the flag protocol were wrong) */
  void delay(size_t di, unsigned long long ns);

  /* sum of curr over the compute regions of all ranks (fp64 accumulation) */
  double checksum();
  /* 64-bit fingerprint of curr over the compute regions of all ranks: every cell's bit pattern is mixed with its
     global coordinate and the results are added modulo 2^64, so the value does not depend on the partition or on the
     order of summation, and two runs agree exactly when (up to a 2^-64 accident) every cell agrees bit for bit */
  unsigned long long fingerprint();
  /* value at global coordinate p if this rank owns it (returns false otherwise) */
  bool probe(const Dim3 &p, float *out);

private:
  struct Impl;
  void build(size_t x, size_t y, size_t z, Comm *comm);
  void fill(bool random, unsigned seed);
  Options opt_;
  std::unique_ptr<DistributedDomain> dd_;
  DataHandle<float> h_;
  std::unique_ptr<Impl> impl_;
  std::atomic<unsigned long long> launches_{0};
};

namespace detail {
/* One entry of a kernel's tile table as a box of cells (compute coordinates of the subdomain, clipped to the region the
   table covers). faces / dirs: the subdomain faces (Jacobi: bit 0..5 = -x +x -y +y -z +z) or push directions (Astaroth:
   bit 0..7 = -y +y -z +z and the four yz diagonals) the tile pushes through. shape: 0 wide, 1 narrow. */
struct TileBox {
  Dim3 lo, hi;
  unsigned faces;
  int shape, field;
};
/* test hooks (host only, no GPU): the table the Jacobi TMA kernel walks for region [lo,hi) of a subdomain of size sz —
   in launch order — and the number of tiles counted per face of countMask */
std::vector<TileBox> jacobi_tile_table(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int tileZ, unsigned faceMask, unsigned countMask, bool narrow,
                                       bool flaggedFirst, unsigned nFace[6], int tailPlanes = 0, int tailDepth = 8);
} // namespace detail

/* test hook: how many of `values` does the kernel's 3-instruction x/6 round differently from IEEE division (expect 0) */
unsigned jacobi_div6_mismatches(const std::vector<float> &values);

/* the value init_random() gives global cell p of a domain of size `global` */
float jacobi_random_value(unsigned seed, const Dim3 &p, const Dim3 &global);

/* host reference of one Jacobi iteration on a dense periodic X*Y*Z array (same arithmetic and order as the kernel) */
void jacobi_host_step(std::vector<float> &next, const std::vector<float> &curr, const Dim3 &size);

} // namespace stencil
