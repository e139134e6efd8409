/* strong: exchange-only strong scaling — radius 3 on all 26 directions, four fp32 quantities, nIters x (exchange(); swap()).
   CLI and the CSV line follow the reference's bin/strong.cu (positional X Y Z nIters; --kernel --peer --colo --staged
   --naive); the positional size is the whole domain. Output on rank 0 (the reference prints it only when
   built with EXCHANGE_STATS; here the timers are switched on at run time):
     weak,<methods>,x,y,z,x*y*z,<B staged>,<B colo>,<B peer>,<B direct>,iters,gpus,nodes,ranks,mpi_topo,node_gpus,peer_en,
          placement,realize,plan,create,exchange,swap
   (the reference's strong.cu labels its row "weak" too; kept). "B direct" = same-GPU kernel + NvlinkDirect bytes. */
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
  argparse::Parser p("exchange-only strong scaling benchmark");
  p.no_unrecognized();
  p.add_positional(x)->required();
  p.add_positional(y)->required();
  p.add_positional(z)->required();
  p.add_positional(nIters)->required();
  margs.add_to(p);
  p.add_flag(naive, "--naive")->help("trivial placement instead of node-aware");
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
    dd.set_radius(3);
    dd.set_placement(naive ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware);
    dd.set_exchange_stats(true);
    for (const char *name : {"d0", "d1", "d2", "d3"}) dd.add_data<float>(name);
    dd.realize();
    comm.barrier();
    for (int iter = 0; iter < nIters; ++iter) {
      dd.exchange();
      dd.swap();
    }
    comm.barrier();
    if (job.rank == 0)
      std::printf("weak,%s,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%lu,%d,%d,%d,%d,%e,%e,%e,%e,%e,%e,%e,%e,%e\n", method_label(methods).c_str(), x, y, z,
                  x * y * z, dd.exchange_bytes_for_method(MethodFlags::CudaMpi), dd.exchange_bytes_for_method(MethodFlags::CudaMpiColocated),
                  dd.exchange_bytes_for_method(MethodFlags::CudaMemcpyPeer),
                  dd.exchange_bytes_for_method(MethodFlags::CudaKernel | MethodFlags::NvlinkDirect), nIters, job.numGpus, job.numNodes,
                  job.size, dd.timeMpiTopo_, dd.timeNodeGpus_, dd.timePeerEn_, dd.timePlacement_, dd.timeRealize_, dd.timePlan_,
                  dd.timeCreate_, dd.timeExchange_, dd.timeSwap_);
  }
  stencil::finalize_world();
  return 0;
}
