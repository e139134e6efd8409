/* jacobi3d: 7-point Jacobi heat solver, weak-scaled by the number of subdomains.
   CLI and CSV line follow the reference's bin/jacobi3d.cu:110-128,378-379:
     jacobi3d X Y Z [--iters/-n N] [--staged --colo --peer --kernel --direct --cuda-aware-mpi] [--trivial] [--no-overlap]
              [--paraview --prefix P --period/-q Q]
   prints on rank 0:  jacobi3d,<methods>,<ranks>,<devCount>,<x>,<y>,<z>,<min iter s>,<trimean iter s>
   Additions: --direct (MethodFlags::NvlinkDirect), --no-fused (use the reference's exchange() schedule instead of
   the fused compute+push kernel), --exact-size (do not weak-scale the positional size), --tile-z. */
#include <cmath>
#include <cstdio>
#include <iostream>

#include "common.hpp"
#include "stencil/jacobi.hpp"
#include "stencil/statistics.hpp"

int main(int argc, char **argv) {
  MethodArgs methodArgs;
  bool trivial = false, noOverlap = false, paraview = false, noFused = false, exactSize = false, ldg = false, naturalOrder = false;
  bool deviceTime = false;
  int verifyIters = 0, delayRank = 0;
  double delayUs = 0;
  size_t x = 512, y = 512, z = 512;
  std::string prefix;
  int iters = 30, period = -1, tileZ = 0;

  argparse::Parser parser("7-point Jacobi on a periodic box with a hot and a cold sphere");
  methodArgs.add_to(parser);
  parser.add_flag(trivial, "--trivial")->help("skip node-aware placement");
  parser.add_flag(noOverlap, "--no-overlap")->help("don't overlap communication and computation");
  parser.add_flag(noFused, "--no-fused")->help("use exchange() + separate kernels instead of the fused kernel");
  parser.add_flag(ldg, "--ldg")->help("use the plain-load kernel variant instead of the TMA pipeline");
  parser.add_flag(naturalOrder, "--natural-order")->help("do not schedule the tiles that push to other GPUs first");
  parser.add_flag(deviceTime, "--device-time")->help("also print the free-running device time per iteration (no host sync in the loop)");
  parser.add_option(verifyIters, "--verify")->help("instead of timing: N iterations from a random field with the plain kernel + exchange() and "
                                                    "with the fused kernel, then compare the fingerprints of the two fields");
  parser.add_option(delayUs, "--delay-us")->help("--verify: hold back one GPU by this many microseconds before every fused iteration");
  parser.add_option(delayRank, "--delay-rank")->help("--verify: the rank whose first GPU is held back (default 0)");
  parser.add_flag(exactSize, "--exact-size")->help("use X Y Z as the whole domain (no weak scaling)");
  parser.add_option(prefix, "--prefix")->help("prefix for paraview files");
  parser.add_flag(paraview, "--paraview")->help("dump paraview files");
  parser.add_option(iters, "--iters", "-n")->help("number of iterations");
  parser.add_option(period, "--period", "-q")->help("iterations between checkpoints");
  parser.add_option(tileZ, "--tile-z")->help("planes marched per CTA (default: chosen from the subdomain size)");
  parser.add_positional(x)->required();
  parser.add_positional(y)->required();
  parser.add_positional(z)->required();
  if (!parser.parse(argc, argv)) {
    std::cerr << parser.help() << "\n";
    return EXIT_FAILURE;
  }
  if (parser.need_help()) {
    std::cerr << parser.help() << "\n";
    return 0;
  }
  if (period <= 0) period = std::max(1, iters / 10);

  stencil::Comm &comm = stencil::world();
  const int rank = comm.rank(), size = comm.size();
  int devCount = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&devCount));
  int numSubdoms;
  {
    MpiTopology topo(comm);
    numSubdoms = size / topo.colocated_size() * devCount;
  }
  if (rank == 0) std::cerr << "assuming " << numSubdoms << " subdomains\n";
  if (!exactSize) {
    const double s = std::pow(double(numSubdoms), 0.33333);
    x = size_t(double(x) * s + 0.5);
    y = size_t(double(y) * s + 0.5);
    z = size_t(double(z) * s + 0.5);
  }

  const MethodFlags methods = methodArgs.flags();

  if (verifyIters > 0) {
    /* fused == plain, bit for bit, on this size and these GPUs; the fingerprint does not depend on the partition */
    unsigned long long fp[2] = {0, 0};
    for (int pass = 0; pass < 2; ++pass) {
      stencil::Jacobi3D::Options opt;
      opt.methods = methods;
      opt.placement = trivial ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware;
      opt.mode = pass == 0 ? stencil::Jacobi3D::Mode::NoOverlap : stencil::Jacobi3D::Mode::Fused;
      opt.tileZ = tileZ;
      opt.useTma = !ldg;
      opt.flaggedFirst = !naturalOrder;
      stencil::Jacobi3D solver(x, y, z, opt);
      solver.init_random(12345);
      for (int iter = 0; iter < verifyIters; ++iter) {
        if (pass == 1 && delayUs > 0 && rank == delayRank) solver.delay(0, (unsigned long long)(delayUs * 1e3));
        solver.step();
      }
      fp[pass] = solver.fingerprint();
      if (pass == 1 && solver.mode() != stencil::Jacobi3D::Mode::Fused && rank == 0) std::cerr << "fused mode not available: compared plain with overlap\n";
    }
    const bool ok = fp[0] == fp[1];
    if (rank == 0) {
      char line[256];
      std::snprintf(line, sizeof(line), "jacobi3d-verify,%d,%d,%zu,%zu,%zu,%d,%016llx,%016llx,%s", size, devCount, x, y, z, verifyIters, fp[0], fp[1],
                    ok ? "ok" : "FAILED");
      std::cout << line << "\n";
    }
    stencil::finalize_world();
    return ok ? 0 : 1;
  }

  Statistics iterTime;
  {
    stencil::Jacobi3D::Options opt;
    opt.methods = methods;
    opt.placement = trivial ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware;
    opt.mode = noOverlap ? stencil::Jacobi3D::Mode::NoOverlap : (noFused ? stencil::Jacobi3D::Mode::Overlap : stencil::Jacobi3D::Mode::Fused);
    opt.tileZ = tileZ;
    opt.useTma = !ldg;
    opt.flaggedFirst = !naturalOrder;
    opt.planFile = true;
    stencil::Jacobi3D solver(x, y, z, opt);
    comm.barrier();
    if (paraview) solver.domain().write_paraview(prefix + "jacobi3d_init");
    for (int iter = 0; iter < iters; ++iter) {
      /* per-iteration time as the reference measures it: wall clock around launch .. completion, max over ranks */
      const double t0 = stencil::Comm::wtime();
      solver.step();
      solver.sync();
      iterTime.insert(comm.allreduce_max(stencil::Comm::wtime() - t0));
      if (paraview && iter % period == 0) solver.domain().write_paraview(prefix + "jacobi3d_" + std::to_string(iter));
    }
    if (paraview) solver.domain().write_paraview(prefix + "jacobi3d_final");
    if (deviceTime) {
      const double secs = solver.run(iters);
      if (rank == 0) std::cout << "jacobi3d-device," << size << "," << devCount << "," << x << "," << y << "," << z << "," << secs / iters << "\n";
    }
  }
  if (rank == 0)
    std::cout << "jacobi3d," << method_string(methods) << "," << size << "," << devCount << "," << x << "," << y << "," << z << ","
              << iterTime.min() << "," << iterTime.trimean() << "\n";
  stencil::finalize_world();
  return 0;
}
