/* strong: exchange-only strong scaling — radius 3 on all 26 directions, four fp32 quantities, nIters x (exchange(); swap()).
   CLI and the CSV line follow the reference's bin/strong.cu (positional X Y Z nIters; --kernel --peer --colo --staged
   --naive); the positional size is the whole domain. Output on rank 0 (the reference prints it only when
   built with EXCHANGE_STATS; here the timers are switched on at run time):
     weak,<methods>,x,y,z,x*y*z,<B staged>,<B colo>,<B peer>,<B direct>,iters,gpus,nodes,ranks,mpi_topo,node_gpus,peer_en,
          placement,realize,plan,create,exchange,swap
   (the reference's strong.cu labels its row "weak" too; kept). "B direct" = same-GPU kernel + NvlinkDirect bytes.
   Additions for BASELINE's strong-scaling config (uneven radius, overlapped compute):
     --xplus A --xminus B   radius A towards +x and B towards -x on every direction that has an x component (3 elsewhere)
     --compute              every iteration also advances quantity d0 with a user-style stencil kernel (mean of the cells
                            within the radius along x and of the y / z face neighbours), interior overlapped with
                            exchange_async(), exterior slabs after it, no host wait when the plan is device-only;
                            prints a second line  strong-compute,<iters>,<device s per iteration>,<Gcell/s>  */
#include <algorithm>
#include <cstdio>
#include <iostream>
#include <vector>

#include "common.hpp"

/* next = mean of curr over x-rm .. x+rp (without x itself) and the four y / z face neighbours: reads exactly the halo
   the radius asks for on the x faces */
__global__ void relax_kernel(Accessor<float> dst, Accessor<float> src, Rect3 box, int rp, int rm) {
  for (int64_t z = box.lo.z + blockIdx.z; z < box.hi.z; z += gridDim.z)
    for (int64_t y = box.lo.y + blockIdx.y * blockDim.y + threadIdx.y; y < box.hi.y; y += gridDim.y * blockDim.y)
      for (int64_t x = box.lo.x + blockIdx.x * blockDim.x + threadIdx.x; x < box.hi.x; x += gridDim.x * blockDim.x) {
        const Dim3 p(x, y, z);
        float v = src[p + Dim3(0, 1, 0)] + src[p - Dim3(0, 1, 0)] + src[p + Dim3(0, 0, 1)] + src[p - Dim3(0, 0, 1)];
        for (int k = 1; k <= rp; ++k) v += src[p + Dim3(k, 0, 0)];
        for (int k = 1; k <= rm; ++k) v += src[p - Dim3(k, 0, 0)];
        dst[p] = v / float(4 + rp + rm);
      }
}

int main(int argc, char **argv) {
  stencil::Comm &comm = stencil::world();
  const JobShape job = job_shape(comm);
  size_t x = 512, y = 512, z = 512;
  int nIters = 30;
  bool naive = false, compute = false;
  int xplus = 3, xminus = 3;
  MethodArgs margs;
  argparse::Parser p("exchange-only strong scaling benchmark");
  p.no_unrecognized();
  p.add_positional(x)->required();
  p.add_positional(y)->required();
  p.add_positional(z)->required();
  p.add_positional(nIters)->required();
  margs.add_to(p);
  p.add_flag(naive, "--naive")->help("trivial placement instead of node-aware");
  p.add_flag(compute, "--compute")->help("also advance d0 with a stencil kernel, interior overlapped with the exchange");
  p.add_option(xplus, "--xplus")->help("radius towards +x (default 3)");
  p.add_option(xminus, "--xminus")->help("radius towards -x (default 3)");
  if (!p.parse(argc, argv)) {
    std::cout << p.help() << "\n";
    return EXIT_FAILURE;
  }
  if (p.need_help()) {
    std::cout << p.help() << "\n";
    return 0;
  }

  const MethodFlags methods = margs.flags();
  if (job.rank == 0) {
    std::cerr << job.numGpus << " subdomains " << job.size << " ranks: " << x << "," << y << "," << z << "=" << x * y * z << "\n";
    std::cerr << (naive ? "naive placement\n" : "node-aware placement\n");
    std::cerr << "methods: " << method_label(methods) << "\n";
  }
  {
    DistributedDomain dd(x, y, z);
    dd.set_methods(methods);
    Radius radius = Radius::constant(3);
    if (xplus != 3 || xminus != 3)
      for (int dz = -1; dz <= 1; ++dz)
        for (int dy = -1; dy <= 1; ++dy) {
          radius.dir(1, dy, dz) = size_t(xplus);
          radius.dir(-1, dy, dz) = size_t(xminus);
        }
    dd.set_radius(radius);
    dd.set_placement(naive ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware);
    dd.set_exchange_stats(!compute);
    std::vector<DataHandle<float>> handles;
    for (const char *name : {"d0", "d1", "d2", "d3"}) handles.push_back(dd.add_data<float>(name));
    dd.realize();
    comm.barrier();
    double computeSeconds = 0;
    if (!compute) {
      for (int iter = 0; iter < nIters; ++iter) {
        dd.exchange();
        dd.swap();
      }
    } else {
      std::vector<LocalDomain> &doms = dd.domains();
      const std::vector<Rect3> interior = dd.get_interior();
      const std::vector<std::vector<Rect3>> exterior = dd.get_exterior();
      const bool hostFree = dd.exchange_device_only();
      const dim3 block(32, 8, 1), grid(8, 16, 32);
      std::vector<RcStream> streams;
      std::vector<cudaEvent_t> done(doms.size()), t0(doms.size()), t1(doms.size());
      for (size_t i = 0; i < doms.size(); ++i) {
        CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
        streams.emplace_back(doms[i].gpu());
        CUDA_RUNTIME(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
        CUDA_RUNTIME(cudaEventCreate(&t0[i]));
        CUDA_RUNTIME(cudaEventCreate(&t1[i]));
      }
      auto iterate = [&](int n, bool first) {
        for (int it = 0; it < n; ++it) {
          if (!(first && it == 0))
            for (size_t i = 0; i < doms.size(); ++i) dd.exchange_after(i, done[i]);
          dd.exchange_async();
          for (size_t i = 0; i < doms.size(); ++i) {
            CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
            relax_kernel<<<grid, block, 0, streams[i]>>>(doms[i].get_next_accessor(handles[0]), doms[i].get_curr_accessor(handles[0]), interior[i], xplus, xminus);
          }
          if (!hostFree) dd.exchange_wait();
          for (size_t i = 0; i < doms.size(); ++i) {
            CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
            if (hostFree) CUDA_RUNTIME(cudaStreamWaitEvent(streams[i], dd.exchange_done_event(i), 0));
            for (const Rect3 &slab : exterior[i])
              if (slab.extent().flatten() > 0)
                relax_kernel<<<grid, block, 0, streams[i]>>>(doms[i].get_next_accessor(handles[0]), doms[i].get_curr_accessor(handles[0]), slab, xplus, xminus);
            CUDA_RUNTIME(cudaEventRecord(done[i], streams[i]));
          }
          dd.swap();
        }
      };
      auto drain = [&]() {
        for (size_t i = 0; i < doms.size(); ++i) {
          CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
          CUDA_RUNTIME(cudaStreamSynchronize(streams[i]));
        }
        dd.exchange_wait();
      };
      iterate(2, true); // warm up
      drain();
      comm.barrier();
      for (size_t i = 0; i < doms.size(); ++i) {
        CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
        CUDA_RUNTIME(cudaEventRecord(t0[i], streams[i]));
      }
      iterate(nIters, false);
      float worst = 0;
      for (size_t i = 0; i < doms.size(); ++i) {
        CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
        CUDA_RUNTIME(cudaEventRecord(t1[i], streams[i]));
      }
      drain();
      for (size_t i = 0; i < doms.size(); ++i) {
        float ms = 0;
        CUDA_RUNTIME(cudaEventElapsedTime(&ms, t0[i], t1[i]));
        worst = std::max(worst, ms);
        cudaEventDestroy(done[i]);
        cudaEventDestroy(t0[i]);
        cudaEventDestroy(t1[i]);
      }
      computeSeconds = comm.allreduce_max(double(worst) * 1e-3);
    }
    comm.barrier();
    if (job.rank == 0)
      std::printf("weak,%s,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%d,%d,%d,%d,%e,%e,%e,%e,%e,%e,%e,%e,%e\n", method_label(methods).c_str(), x, y, z,
                  x * y * z, dd.exchange_bytes_for_method(MethodFlags::CudaMpi), dd.exchange_bytes_for_method(MethodFlags::CudaMpiColocated),
                  dd.exchange_bytes_for_method(MethodFlags::CudaMemcpyPeer),
                  dd.exchange_bytes_for_method(MethodFlags::CudaKernel | MethodFlags::NvlinkDirect), nIters, job.numGpus, job.numNodes,
                  job.size, dd.timeMpiTopo_, dd.timeNodeGpus_, dd.timePeerEn_, dd.timePlacement_, dd.timeRealize_, dd.timePlan_,
                  dd.timeCreate_, dd.timeExchange_, dd.timeSwap_);
    if (job.rank == 0 && compute)
      std::printf("strong-compute,%d,%e,%f\n", nIters, computeSeconds / nIters, double(x) * double(y) * double(z) / (computeSeconds / nIters) / 1e9);
  }
  stencil::finalize_world();
  return 0;
}
