/* astaroth-sim: radius-3, multi-field stencil workload with interior/exterior overlap.
   CLI follows the reference's bin/astaroth_sim.cu:87-101 (--x --y --z default 512 = size per run, --remote
   --cuda-aware-mpi --colocated --peer --kernel --trivial, -f/--file accepted and unused). The reference runs exactly
   one fp32 field, 5 iterations and prints no timing; defaults here reproduce that workload, and the options
   --fields N, --double, --star (real 19-point radius-3 stencil), --cross (that stencil plus cross-derivative terms), --iters, --no-overlap, --weak (scale the size by
   numGpus^0.33333 like the other apps) cover BASELINE's "8 fp64 fields, radius 3, 256^3 per GPU" config.
   Adds one CSV line on rank 0:
     astaroth-sim,<methods>,<ranks>,<devCount>,x,y,z,<fields>,<fp32|fp64>,<avg7|star19>,<iters>,<s per iter>,<exchange s per iter>,<Gcell/s> */
#include <algorithm>
#include <cmath>
#include <iostream>

#include "common.hpp"
#include "stencil/astaroth.hpp"

int main(int argc, char **argv) {
  size_t x = 512, y = 512, z = 512;
  int iters = 5, fields = 1, tileZ = 0;
  bool remote = false, cudaAware = false, colocated = false, peer = false, kernel = false, direct = false, trivial = false;
  bool fp64 = false, star = false, cross = false, noOverlap = false, forceOverlap = false, weak = false, check = false;
  std::string file;
  argparse::Parser p("radius-3 multi-field stencil (Astaroth stand-in)");
  p.add_option(x, "--x");
  p.add_option(y, "--y");
  p.add_option(z, "--z");
  p.add_flag(remote, "--remote")->help("host-staged transport");
  p.add_flag(cudaAware, "--cuda-aware-mpi")->help("accepted for compatibility; needs a CUDA-aware MPI Comm");
  p.add_flag(colocated, "--colocated")->help("colocated CUDA IPC transport");
  p.add_flag(peer, "--peer")->help("pack + cudaMemcpyPeer");
  p.add_flag(kernel, "--kernel")->help("same-GPU copies");
  p.add_flag(direct, "--direct")->help("device-initiated NVLink stores");
  p.add_flag(trivial, "--trivial")->help("skip node-aware placement");
  p.add_option(file, "--file", "-f")->help("unused (reference compatibility)");
  p.add_option(iters, "--iters");
  p.add_option(fields, "--fields");
  p.add_option(tileZ, "--tile-z")->help("planes marched per CTA (default 48, balanced over the subdomain)");
  p.add_flag(fp64, "--double");
  p.add_flag(star, "--star");
  p.add_flag(star, "--star19")->help("same as --star (torchrun's parser mistakes --star for an abbreviation of its --start-method)");
  p.add_flag(cross, "--cross")->help("19-point stencil plus cross-derivative terms: reads edge and corner halos too");
  p.add_flag(noOverlap, "--no-overlap");
  p.add_flag(forceOverlap, "--overlap")->help("keep the interior/exterior schedule even where the engine would drop it");
  p.add_flag(weak, "--weak");
  p.add_flag(check, "--check")->help("small domains: compare every rank's cells of field 0 and of the last field with a host loop");
  if (!p.parse(argc, argv)) {
    std::cerr << p.help() << "\n";
    return EXIT_FAILURE;
  }
  if (p.need_help()) {
    std::cerr << p.help() << "\n";
    return 0;
  }
  stencil::Comm &comm = stencil::world();
  const JobShape job = job_shape(comm);
  if (job.rank == 0) std::cerr << "assuming " << job.numGpus << " subdomains\n";
  if (weak) {
    x = weak_scale(x, job.numGpus);
    y = weak_scale(y, job.numGpus);
    z = weak_scale(z, job.numGpus);
  }
  if (job.rank == 0) std::cerr << "domain: " << x << "," << y << "," << z << "\n";
  MethodFlags methods = MethodFlags::None;
  if (remote) methods |= MethodFlags::CudaMpi;
  if (colocated) methods |= MethodFlags::CudaMpiColocated;
  if (peer) methods |= MethodFlags::CudaMemcpyPeer;
  if (kernel) methods |= MethodFlags::CudaKernel;
  if (direct) methods |= MethodFlags::NvlinkDirect;
  if (methods == MethodFlags::None) methods = MethodFlags::All;
  double perIter = 0, exPerIter = 0;
  bool checkFailed = false;
  {
    stencil::AstarothSim::Options opt;
    opt.methods = methods;
    opt.placement = trivial ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware;
    opt.fields = fields;
    opt.fp64 = fp64;
    opt.stencil = cross ? stencil::AstarothSim::Stencil::Star19Cross : (star ? stencil::AstarothSim::Stencil::Star19 : stencil::AstarothSim::Stencil::Average7);
    opt.overlap = !noOverlap;
    opt.autoSchedule = !forceOverlap;
    opt.planFile = true;
    if (tileZ > 0) opt.tileZ = tileZ;
    stencil::AstarothSim sim(x, y, z, opt);
    if (check) {
      /* every rank advances the whole (small) domain on the host from the device's own initial field and compares
         the cells it owns; ranks agree on the initial field because it is a function of the global coordinate */
      const Dim3 size(static_cast<int64_t>(x), static_cast<int64_t>(y), static_cast<int64_t>(z));
      double worst = 0;
      for (int f : {0, fields - 1}) {
        std::vector<double> ref(size_t(size.flatten())), tmp;
        for (int64_t k = 0; k < size.z; ++k)
          for (int64_t j = 0; j < size.y; ++j)
            for (int64_t i = 0; i < size.x; ++i) ref[size_t((k * size.y + j) * size.x + i)] = stencil::astaroth_init_value(Dim3(i, j, k), opt.period, fp64);
        sim.init();
        const std::vector<double> first = sim.gather(f);
        for (size_t i = 0; i < first.size(); ++i)
          if (first[i] == first[i]) ref[i] = first[i]; // my cells: bit-identical start (device sin differs in the last ulp)
        for (int it = 0; it < iters; ++it) {
          sim.step();
          stencil::astaroth_host_step(tmp, ref, size, opt.stencil, fp64, opt.dt, opt.crossWeight);
          ref.swap(tmp);
        }
        const std::vector<double> got = sim.gather(f);
        for (size_t i = 0; i < got.size(); ++i)
          if (got[i] == got[i]) worst = std::max(worst, std::abs(got[i] - ref[i]));
      }
      worst = comm.allreduce_max(worst);
      const double tol = fp64 ? 1e-11 : 1e-4; // cells next to another rank start from that rank's sin(): ulp-level differences grow
      if (job.rank == 0) std::cout << "astaroth-check," << job.size << "," << x << "," << y << "," << z << "," << worst << "," << (worst < tol ? "ok" : "FAILED") << "\n";
      checkFailed = !(worst < tol);
    } else {
    sim.run(1); // warm up
    double ex = 0;
    const double t = sim.run(iters, &ex);
    perIter = t / iters;
    exPerIter = ex / iters;
    }
  } // the engine goes away before the process group does
  if (check) {
    stencil::finalize_world();
    return checkFailed ? 1 : 0;
  }
  if (job.rank == 0)
    std::cout << "astaroth-sim," << method_string(methods) << "," << job.size << "," << job.devCount << "," << x << "," << y << "," << z << ","
              << fields << "," << (fp64 ? "fp64" : "fp32") << "," << (cross ? "star19cross" : (star ? "star19" : "avg7")) << "," << iters << "," << perIter << ","
              << exPerIter << "," << double(x) * double(y) * double(z) / perIter / 1e9 << "\n";
  stencil::finalize_world();
  return 0;
}
