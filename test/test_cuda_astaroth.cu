/* AstarothSim against the host loop: both stencils, fp32 and fp64, several fields, several subdomains, overlap on/off.
   Average7 in fp32 is the reference's workload and must match bit for bit; Star19 is compared with an absolute tolerance of a
   few ulp of 1 because the device contracts a*b+c into FMAs. */
#include "harness.hpp"
#include <cmath>
#include <vector>
#include "stencil/astaroth.hpp"

namespace {
double max_abs_err(const Dim3 &size, stencil::AstarothSim::Stencil st, bool fp64, int fields, const std::vector<int> &gpus, bool overlap, int iters,
                   MethodFlags methods = MethodFlags::All, bool useTma = true, bool narrowTiles = true, int tailPlanes = 0) {
  stencil::AstarothSim::Options opt;
  opt.stencil = st;
  opt.fp64 = fp64;
  opt.fields = fields;
  opt.gpus = gpus;
  opt.overlap = overlap;
  opt.autoSchedule = false; // test exactly the schedule asked for
  opt.tileZ = 8;
  opt.methods = methods;
  opt.useTma = useTma;
  opt.narrowTiles = narrowTiles;
  opt.tailPlanes = tailPlanes;
  opt.tailDepth = 4;
  stencil::AstarothSim sim(size_t(size.x), size_t(size.y), size_t(size.z), opt);
  /* start from the device's own initial field: the device and host sin() differ in the last bits of a double */
  std::vector<double> ref = sim.gather(0), tmp;
  double initErr = 0;
  for (int64_t z = 0; z < size.z; ++z)
    for (int64_t y = 0; y < size.y; ++y)
      for (int64_t x = 0; x < size.x; ++x)
        initErr = std::max(initErr, std::fabs(ref[size_t((z * size.y + y) * size.x + x)] - stencil::astaroth_init_value(Dim3(x, y, z), opt.period, fp64)));
  if (!(initErr < 1e-6)) return 1e30; // the initial condition itself is wrong
  for (int it = 0; it < iters; ++it) {
    sim.step();
    stencil::astaroth_host_step(tmp, ref, size, st, fp64, opt.dt, opt.crossWeight);
    ref.swap(tmp);
  }
  double worst = 0;
  for (int f = 0; f < fields; ++f) {
    const std::vector<double> got = sim.gather(f);
    for (size_t i = 0; i < got.size(); ++i) {
      const double e = std::fabs(got[i] - ref[i]); // field values are O(1)
      if (!(e <= worst)) worst = e; // also catches NaN
    }
  }
  return worst;
}
} // namespace

TEST_CASE("astaroth: reference workload (1 x fp32, 7-point average) is bit exact", "[cuda][astaroth]") {
  using S = stencil::AstarothSim::Stencil;
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Average7, false, 1, {0}, true, 4) == 0);
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Average7, false, 1, {0, 0}, true, 4) == 0);
  REQUIRE(max_abs_err(Dim3(41, 23, 19), S::Average7, false, 2, {0, 0, 0}, false, 3) == 0);
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Average7, true, 2, {0, 0}, true, 3) == 0);
}

TEST_CASE("astaroth: 19-point radius-3 stencil, fp64 and fp32, many fields and subdomains", "[cuda][astaroth]") {
  using S = stencil::AstarothSim::Stencil;
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Star19, true, 1, {0}, true, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(44, 26, 22), S::Star19, true, 8, {0, 0}, true, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(45, 27, 23), S::Star19, true, 3, {0, 0, 0, 0}, false, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(48, 26, 22), S::Star19, false, 4, {0, 0}, true, 3) < 5e-6);
  REQUIRE(max_abs_err(Dim3(44, 26, 22), S::Star19, true, 2, {0, 0}, true, 3, MethodFlags::CudaMpi) < 1e-13);
  /* wide enough for the TMA kernel in both precisions (tile = 128 fp32 / 64 fp64 cells), partial tiles in x, y and z */
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19, false, 2, {0}, true, 3) < 5e-6);
  REQUIRE(max_abs_err(Dim3(300, 40, 19), S::Star19, false, 3, {0, 0}, true, 3) < 5e-6);
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19, true, 2, {0, 0}, true, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19, true, 2, {0, 0}, false, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(260, 33, 18), S::Star19, false, 2, {0}, false, 3) < 5e-6);
  /* subdomains that wrap onto themselves in x: from the second step on the kernel writes the x halos and exchange() skips them */
  REQUIRE(max_abs_err(Dim3(136, 24, 20), S::Star19, true, 2, {0}, false, 5) < 1e-13);
  REQUIRE(max_abs_err(Dim3(64, 40, 64), S::Star19, true, 3, {0, 0}, false, 5) < 1e-13);
  REQUIRE(max_abs_err(Dim3(128, 40, 64), S::Star19, false, 3, {0, 0, 0, 0}, false, 5) < 2e-5);
  REQUIRE(max_abs_err(Dim3(132, 24, 20), S::Star19, false, 1, {0}, false, 4) < 2e-5);
  /* row lengths that are not a multiple of the vector: element stores into the opposite halo */
  REQUIRE(max_abs_err(Dim3(133, 24, 20), S::Star19, true, 2, {0}, false, 5) < 1e-13);
  REQUIRE(max_abs_err(Dim3(131, 24, 20), S::Star19, false, 2, {0}, false, 5) < 2e-5);
  REQUIRE(max_abs_err(Dim3(70, 21, 40), S::Star19, false, 2, {0, 0}, false, 5) < 2e-5);
  /* several subdomains on one GPU: fused y / z pushes between them ordered by events, 8 steps */
  REQUIRE(max_abs_err(Dim3(64, 40, 64), S::Star19, true, 2, {0, 0, 0, 0}, false, 8) < 1e-13);
  REQUIRE(max_abs_err(Dim3(64, 35, 50), S::Star19, true, 2, {0, 0, 0}, false, 8) < 1e-13);
}

TEST_CASE("astaroth: rows that are not a multiple of the tile use narrow remainder tiles", "[cuda][astaroth][narrow]") {
  using S = stencil::AstarothSim::Stencil;
  /* fp64: wide 64 x 16, narrow 16 x 64; fp32: wide 128 x 16, narrow 32 x 64 */
  REQUIRE(max_abs_err(Dim3(70, 70, 20), S::Star19, true, 2, {0}, false, 5) < 1e-13);     // 64 + 6: one narrow column, two narrow row blocks
  REQUIRE(max_abs_err(Dim3(100, 40, 20), S::Star19, true, 2, {0, 0}, false, 5) < 1e-13); // 64 + 36: three narrow columns
  REQUIRE(max_abs_err(Dim3(120, 24, 20), S::Star19, true, 1, {0}, false, 4) < 1e-13);    // 64 + 56: a second wide column instead
  REQUIRE(max_abs_err(Dim3(200, 70, 14), S::Star19, false, 2, {0}, false, 5) < 2e-5);    // 128 + 72: three narrow fp32 columns
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19, true, 2, {0, 0}, true, 3) < 1e-13);  // overlap schedule: interior / exterior tile tables
  REQUIRE(max_abs_err(Dim3(70, 70, 20), S::Star19, true, 2, {0}, false, 5, MethodFlags::All, true, false) < 1e-13); // remainder tiles off
}

TEST_CASE("astaroth: thin z chunks at the top of the subdomain", "[cuda][astaroth][tail]") {
  using S = stencil::AstarothSim::Stencil;
  /* chunk depth 8, tail 8 planes in chunks of 4: needs 24 planes per subdomain */
  REQUIRE(max_abs_err(Dim3(150, 37, 30), S::Star19, true, 2, {0}, false, 4, MethodFlags::All, true, true, 8) < 1e-13);
  REQUIRE(max_abs_err(Dim3(150, 37, 56), S::Star19, true, 2, {0, 0}, false, 5, MethodFlags::All, true, true, 8) < 1e-13); // z pushes by a thin and a deep chunk
  REQUIRE(max_abs_err(Dim3(133, 40, 27), S::Star19, false, 3, {0}, false, 4, MethodFlags::All, true, true, 8) < 2e-5);
  REQUIRE(max_abs_err(Dim3(140, 36, 30), S::Star19Cross, true, 2, {0}, false, 4, MethodFlags::All, true, true, 8) < 1e-12);
}

TEST_CASE("astaroth: cross-derivative stencil reads edge and corner halos", "[cuda][astaroth][cross]") {
  using S = stencil::AstarothSim::Stencil;
  /* plain kernel (one thread per cell), exchange() every step */
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Star19Cross, true, 1, {0}, true, 3, MethodFlags::All, false) < 1e-12);
  REQUIRE(max_abs_err(Dim3(41, 23, 19), S::Star19Cross, true, 2, {0, 0, 0}, false, 3, MethodFlags::All, false) < 1e-12);
  REQUIRE(max_abs_err(Dim3(40, 24, 20), S::Star19Cross, false, 2, {0, 0}, true, 3, MethodFlags::All, false) < 5e-5);
  /* pipelined kernel, overlap schedule (exchange() every step, tile tables of interior and exterior) */
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19Cross, true, 2, {0, 0}, true, 3) < 1e-12);
  REQUIRE(max_abs_err(Dim3(150, 37, 21), S::Star19Cross, false, 2, {0}, true, 3) < 5e-5);
  /* pipelined kernel, fused: after the first step every halo — faces, the 12 edges, the 8 corners — comes from the
     kernels' own pushes; a wrong or stale edge / corner block shows up here */
  REQUIRE(max_abs_err(Dim3(64, 40, 64), S::Star19Cross, true, 2, {0}, false, 6) < 1e-12);
  REQUIRE(max_abs_err(Dim3(64, 40, 64), S::Star19Cross, true, 2, {0, 0}, false, 8) < 1e-12);
  REQUIRE(max_abs_err(Dim3(70, 44, 50), S::Star19Cross, true, 3, {0, 0, 0, 0}, false, 8) < 1e-12);
  REQUIRE(max_abs_err(Dim3(64, 35, 50), S::Star19Cross, true, 1, {0, 0, 0}, false, 8) < 1e-12);
  REQUIRE(max_abs_err(Dim3(133, 40, 36), S::Star19Cross, false, 2, {0, 0, 0, 0}, false, 8) < 1e-4);
}

TEST_CASE("astaroth: all visible GPUs", "[cuda][astaroth][multigpu]") {
  int n = 0;
  cudaGetDeviceCount(&n);
  if (n < 2) {
    std::printf("  (1 GPU visible: multi-GPU astaroth not exercised)\n");
    return;
  }
  std::vector<int> gpus;
  for (int i = 0; i < n; ++i) gpus.push_back(i);
  using S = stencil::AstarothSim::Stencil;
  REQUIRE(max_abs_err(Dim3(60, 40, 36), S::Star19, true, 8, gpus, true, 3) < 1e-13);
  REQUIRE(max_abs_err(Dim3(60, 40, 36), S::Average7, false, 1, gpus, true, 3) == 0);
  /* no exchange() after the first step: x halos by the kernel's wrap stores, y / z faces pushed to the other GPUs with
     flag hand-shakes (y and z cuts only: 64-wide rows, 8 iterations so that both buffer parities are reused) */
  REQUIRE(max_abs_err(Dim3(64, 48, 64), S::Star19, true, 3, gpus, false, 8) < 1e-13);
  REQUIRE(max_abs_err(Dim3(128, 52, 44), S::Star19, false, 2, gpus, false, 8) < 5e-5);
  std::vector<int> twice = gpus;
  twice.insert(twice.end(), gpus.begin(), gpus.end());
  REQUIRE(max_abs_err(Dim3(66, 48, 64), S::Star19, true, 2, twice, false, 7) < 1e-13);
  /* all 26 directions from the kernels' pushes across GPUs, checked through the cross terms; odd sizes give uneven
     subdomains and narrow remainder tiles */
  REQUIRE(max_abs_err(Dim3(64, 48, 64), S::Star19Cross, true, 3, gpus, false, 8) < 1e-12);
  REQUIRE(max_abs_err(Dim3(70, 50, 66), S::Star19Cross, true, 2, gpus, false, 9) < 1e-12);
  REQUIRE(max_abs_err(Dim3(70, 50, 66), S::Star19Cross, true, 2, twice, false, 7) < 1e-12);
}
