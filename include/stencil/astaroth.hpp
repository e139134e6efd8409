#pragma once
/* AstarothSim: the multi-field radius-3 workload of the reference's bin/astaroth_sim.cu as a reusable engine.

   Reference behaviour kept (bin/astaroth_sim.cu:14-83, :185-274): radius 3 on all 26 directions, fields initialised
   to sin(2*3.14159/period * (x + y + z)) (global coordinates, period 10) with -10 in the halo, and the schedule
   interior kernel -> exchange() -> exterior slabs -> synchronise -> swap(). The reference ships ONE fp32 field and a
   plain 7-point average as stand-in compute (Stencil::Average7); BASELINE's target config is 8 fp64 fields with a real
   radius-3 stencil, so the engine is templated over the element type, takes any number of fields, and adds
   Stencil::Star19: next = curr + dt * L(curr) with the 6th-order, 19-point radius-3 Laplacian every Astaroth-style
   MHD code is built from. All fields are advanced by ONE launch per region (field index is a grid dimension). */
#include <atomic>
#include <memory>
#include <vector>

#include "stencil/jacobi.hpp"
#include "stencil/stencil.hpp"

namespace stencil {

class AstarothSim {
public:
  /* Star19Cross = Star19 plus cross-derivative terms (6th-order d2/dxdy, d2/dxdz, d2/dydz along the diagonals and a
     d3/dxdydz corner term, weighted by Options::crossWeight): reads all 26 halo directions, as the mixed derivatives of
     a real Astaroth right-hand side do */
  enum class Stencil { Average7, Star19, Star19Cross };
  struct Options {
    MethodFlags methods = MethodFlags::All;
    PlacementStrategy placement = PlacementStrategy::NodeAware;
    Stencil stencil = Stencil::Average7;
    int fields = 1;
    bool fp64 = false;
    bool overlap = true;       // interior kernel, exchange(), exterior slabs (the reference's schedule)
    bool autoSchedule = true;  // let the engine drop the overlap when the TMA kernel is used: thin exterior slabs cannot
                               // fill its tiles, and exchange() + one whole-region launch measured 1.8x faster on B200
    bool freeRunning = true; // without overlap and with device-only transports: exchange_async() + events, the host never waits
    bool fusedXFaces = true; // subdomains that wrap onto themselves in x: the kernel writes the +-x halos, exchange() skips them
    bool fusedYZFaces = true; // ... and pushes its y / z face rows, its yz edge bundles and the x wrap of all of them (xy / xz
                              // edges, corners) into the neighbours' halos (other GPUs over NVLink, flags as in the fused
                              // Jacobi kernel): all 26 directions stay current with no exchange() left in steady state
    bool narrowTiles = true;  // TMA kernel: cover the remainder of rows that are not a multiple of the tile with 8-lane tiles
    bool flaggedFirst = true; // TMA kernel: schedule the tiles that push to ANOTHER GPU first
    bool zigzag = true;       // TMA kernel: neighbouring z chunks march in opposite directions (shared planes hit L2)
    double crossWeight = 0.25; // Star19Cross: weight of the cross-derivative terms
    bool useTma = true; // Star19: TMA + mbarrier pipelined kernel for regions wide enough to fill its tiles
    std::vector<int> gpus;
    double period = 10;
    double dt = 0.01;
    int tileZ = 48; // planes per z chunk asked for (chunks are balanced: 256 planes -> 6 x 43); measured 1 - 2 % over 32 on the weak-scaling shapes
    int tailPlanes = 0; // TMA kernel: the top tailPlanes planes of a subdomain in chunks of tailDepth planes, scheduled last
    int tailDepth = 8;  // (STENCIL_ASTAROTH_TAIL_PLANES / _DEPTH)
    bool planFile = false;
  };

  AstarothSim(size_t x, size_t y, size_t z, const Options &opt);
  ~AstarothSim();

  DistributedDomain &domain() { return *dd_; }
  const Options &options() const { return opt_; }
  void init();
  /* one iteration: interior, exchange, exterior, synchronise, swap */
  void step();
  void sync();
  /* `iters` iterations; returns host wall seconds (max over ranks), in *exchangeSeconds the part spent inside
     exchange(), and in *deviceSeconds the time between CUDA events recorded on every compute stream before the first and
     after the last iteration (max over local GPUs and ranks) */
  double run(int iters, double *exchangeSeconds = nullptr, double *deviceSeconds = nullptr);
  unsigned long long launches() const { return launches_; }
  /* dense host copy [z][y][x] of field f over the whole domain owned by this rank (NaN elsewhere), as double */
  std::vector<double> gather(int field);

private:
  struct Impl;
  void setup_fused_faces();
  Options opt_;
  std::unique_ptr<DistributedDomain> dd_;
  std::unique_ptr<Impl> impl_;
  std::atomic<unsigned long long> launches_{0};
};

namespace detail {
/* test hook (host only, no GPU): the tile table of the Astaroth TMA kernel for region [lo,hi) of a subdomain of size sz,
   in launch order; vecN = cells per 16-byte vector (4 fp32, 2 fp64); TileBox as in stencil/jacobi.hpp */
std::vector<TileBox> astaroth_tile_table(const Dim3 &lo, const Dim3 &hi, const Dim3 &sz, int fields, int tileZ, int vecN, unsigned dirMask,
                                                unsigned flagMask, bool narrow, bool flaggedFirst, unsigned nDir[8], int *tileZUsed, int tailPlanes = 0, int tailDepth = 8);
} // namespace detail

/* host reference: one step of one field on a dense periodic array (same arithmetic order as the kernel) */
void astaroth_host_step(std::vector<double> &next, const std::vector<double> &curr, const Dim3 &size, AstarothSim::Stencil st, bool fp64,
                        double dt, double crossWeight = 0.25);
/* the initial field value at a global coordinate */
double astaroth_init_value(const Dim3 &p, double period, bool fp64);

} // namespace stencil
