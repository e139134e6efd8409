/* weak-exchange: as `weak`, but times the whole loop of nIters exchange() calls (no swap) with one wall clock, max
   over ranks. CLI and CSV line follow the reference's bin/weak_exchange.cu:
     weak,<methods>,x,y,z,x*y*z,<B staged>,<B colo>,<B peer>,<B direct>,iters,gpus,nodes,ranks,<elapsed s>
   Additionally prints the device-timed figure (CUDA events on every exchange stream, max over GPUs and ranks) on a
   second line starting with "weak-exchange-device," so scripts that parse the first line keep working. */
#include <cstdio>
#include <iostream>

#include "common.hpp"

int main(int argc, char **argv) {
  stencil::Comm &comm = stencil::world();
  const JobShape job = job_shape(comm);
  size_t x = 512, y = 512, z = 512;
  int nIters = 30;
  bool naive = false;
  MethodArgs margs;
  argparse::Parser p("exchange-only weak scaling benchmark, loop-timed");
  p.no_unrecognized();
  p.add_positional(x)->required();
  p.add_positional(y)->required();
  p.add_positional(z)->required();
  p.add_positional(nIters)->required();
  margs.add_to(p);
  p.add_flag(naive, "--naive");
  if (!p.parse(argc, argv)) {
    std::cout << p.help() << "\n";
    return EXIT_FAILURE;
  }
  if (p.need_help()) {
    std::cout << p.help() << "\n";
    return 0;
  }
  x = weak_scale(x, job.numGpus);
  y = weak_scale(y, job.numGpus);
  z = weak_scale(z, job.numGpus);
  const MethodFlags methods = margs.flags();
  {
    DistributedDomain dd(x, y, z);
    dd.set_methods(methods);
    dd.set_radius(3);
    dd.set_placement(naive ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware);
    for (const char *name : {"d0", "d1", "d2", "d3"}) dd.add_data<float>(name);
    dd.realize();
    dd.exchange(); // warm up: first-touch of peer mappings
    comm.barrier();
    std::vector<cudaEvent_t> t0(dd.domains().size()), t1(dd.domains().size());
    for (size_t di = 0; di < dd.domains().size(); ++di) {
      CUDA_RUNTIME(cudaSetDevice(dd.domains()[di].gpu()));
      CUDA_RUNTIME(cudaEventCreate(&t0[di]));
      CUDA_RUNTIME(cudaEventCreate(&t1[di]));
      CUDA_RUNTIME(cudaEventRecord(t0[di], dd.exchange_stream(di)));
    }
    double elapsed = stencil::Comm::wtime();
    for (int iter = 0; iter < nIters; ++iter) dd.exchange();
    elapsed = comm.allreduce_max(stencil::Comm::wtime() - elapsed);
    float devMs = 0;
    for (size_t di = 0; di < dd.domains().size(); ++di) {
      CUDA_RUNTIME(cudaSetDevice(dd.domains()[di].gpu()));
      CUDA_RUNTIME(cudaEventRecord(t1[di], dd.exchange_stream(di)));
      CUDA_RUNTIME(cudaEventSynchronize(t1[di]));
      float ms = 0;
      CUDA_RUNTIME(cudaEventElapsedTime(&ms, t0[di], t1[di]));
      devMs = std::max(devMs, ms);
      cudaEventDestroy(t0[di]);
      cudaEventDestroy(t1[di]);
    }
    const double devSeconds = comm.allreduce_max(double(devMs) * 1e-3);
    if (job.rank == 0) {
      std::printf("weak,%s,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%d,%d,%d,%d,%e\n", method_label(methods).c_str(), x, y, z, x * y * z,
                  dd.exchange_bytes_for_method(MethodFlags::CudaMpi), dd.exchange_bytes_for_method(MethodFlags::CudaMpiColocated),
                  dd.exchange_bytes_for_method(MethodFlags::CudaMemcpyPeer),
                  dd.exchange_bytes_for_method(MethodFlags::CudaKernel | MethodFlags::NvlinkDirect), nIters, job.numGpus, job.numNodes,
                  job.size, elapsed);
      std::printf("weak-exchange-device,%s,%d,%e,%e\n", method_label(methods).c_str(), nIters, devSeconds, devSeconds / nIters);
    }
  }
  stencil::finalize_world();
  return 0;
}
