#pragma once
/* shared by the apps: method-selection flags (same spellings as the reference's apps) and the job geometry */
#include <cmath>
#include <string>

#include "stencil/argparse.hpp"
#include "stencil/stencil.hpp"

struct MethodArgs {
  bool staged = false, colo = false, peer = false, kernel = false, direct = false, cudaAware = false;
  void add_to(argparse::Parser &p) {
    p.add_flag(kernel, "--kernel")->help("same-GPU copies (CudaKernel)");
    p.add_flag(peer, "--peer")->help("pack + cudaMemcpyPeer (CudaMemcpyPeer)");
    p.add_flag(colo, "--colo")->help("colocated CUDA IPC transport (CudaMpiColocated)");
    p.add_flag(staged, "--staged")->help("host-staged transport (CudaMpi)");
    p.add_flag(direct, "--direct")->help("device-initiated NVLink stores (NvlinkDirect)");
    /* both spellings of the reference: weak/strong/weak-exchange say --cuda-aware, jacobi3d says --cuda-aware-mpi */
    p.add_flag(cudaAware, "--cuda-aware")->help("device-pointer messages (CudaAwareMpi; needs a CUDA-aware MPI process group)");
    p.add_flag(cudaAware, "--cuda-aware-mpi")->help("same as --cuda-aware");
  }
  MethodFlags flags() const {
    MethodFlags m = MethodFlags::None;
    if (staged) m |= MethodFlags::CudaMpi;
    if (colo) m |= MethodFlags::CudaMpiColocated;
    if (peer) m |= MethodFlags::CudaMemcpyPeer;
    if (kernel) m |= MethodFlags::CudaKernel;
    if (direct) m |= MethodFlags::NvlinkDirect;
    if (cudaAware) m |= MethodFlags::CudaAwareMpi;
    return m == MethodFlags::None ? MethodFlags::All : m;
  }
};

/* label in the reference's style: pieces joined by '/', plus "/all" when every method is on */
inline std::string method_label(MethodFlags m) {
  std::string s = method_string(m);
  if (m == MethodFlags::All) s += "/all";
  return s;
}

struct JobShape {
  int rank, size, devCount, numGpus, numNodes;
};

inline JobShape job_shape(stencil::Comm &comm) {
  JobShape j;
  j.rank = comm.rank();
  j.size = comm.size();
  CUDA_RUNTIME(cudaGetDeviceCount(&j.devCount));
  MpiTopology topo(comm);
  j.numNodes = topo.size() / topo.colocated_size();
  /* subdomains in the job when no GPU is shared (reference: size / colocated_size * devCount) */
  j.numGpus = j.size / topo.colocated_size() * j.devCount;
  return j;
}

/* the reference's weak-scaling rule: round(v * n^0.33333) */
inline size_t weak_scale(size_t v, int n) { return size_t(double(v) * std::pow(double(n), 0.33333) + 0.5); }
