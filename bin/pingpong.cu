/* pingpong: host-memory ping-pong latency/bandwidth between nodes through the process group.
   Reference: bin/pingpong.cu (MPI_Send/Recv between every pair of nodes; positional ranksPerNode; --min --max --iters;
   one line per node pair "hostA-hostB-rpn t0 t1 ..." with the time of 2^n / ranksPerNode byte messages for
   n = min..max). Same CLI and line format over stencil::Comm; --all-ranks additionally measures rank pairs that
   share a node (useful on a single box, where there are no node pairs). */
#include <iostream>
#include <vector>

#include "stencil/argparse.hpp"
#include "stencil/comm.hpp"
#include "stencil/mpi_topology.hpp"

int main(int argc, char **argv) {
  int ranksPerNode = 1, minN = 0, maxN = 27, iters = 10;
  bool allRanks = false;
  argparse::Parser p("process-group ping-pong");
  p.add_positional(ranksPerNode)->required();
  p.add_option(minN, "--min");
  p.add_option(maxN, "--max");
  p.add_option(iters, "--iters");
  p.add_flag(allRanks, "--all-ranks");
  if (!p.parse(argc, argv) || p.need_help()) {
    std::cout << p.help();
    return p.need_help() ? 0 : EXIT_FAILURE;
  }
  stencil::Comm &comm = stencil::world();
  MpiTopology topo(comm);
  const int rank = comm.rank(), size = comm.size();
  std::vector<char> buf(size_t(1) << maxN);
  /* pairs (a < b): first ranks of two different nodes, or every rank pair with --all-ranks */
  for (int a = 0; a < size; ++a)
    for (int b = a + 1; b < size; ++b) {
      const bool nodePair = topo.node_of(a) != topo.node_of(b) && topo.host_of(a) != topo.host_of(b) &&
                            [&] {
                              for (int r = 0; r < a; ++r)
                                if (topo.node_of(r) == topo.node_of(a)) return false;
                              for (int r = 0; r < b; ++r)
                                if (topo.node_of(r) == topo.node_of(b)) return false;
                              return true;
                            }();
      if (!nodePair && !allRanks) continue;
      std::vector<double> times;
      for (int nn = minN; nn <= maxN; ++nn) {
        const size_t bytes = std::max<size_t>(1, (size_t(1) << nn) / size_t(ranksPerNode));
        comm.barrier();
        double t = stencil::Comm::wtime();
        for (int i = 0; i < iters; ++i) {
          if (rank == a) {
            comm.send(b, 77, buf.data(), bytes);
            comm.recv(b, 78, buf.data(), bytes);
          } else if (rank == b) {
            comm.recv(a, 77, buf.data(), bytes);
            comm.send(a, 78, buf.data(), bytes);
          }
        }
        t = (stencil::Comm::wtime() - t) / iters;
        times.push_back(comm.allreduce_max(rank == a ? t : 0.0));
      }
      if (rank == 0) {
        std::cout << topo.host_of(a) << "-" << topo.host_of(b) << "-" << ranksPerNode;
        for (double t : times) std::cout << " " << t;
        std::cout << "\n";
      }
    }
  stencil::finalize_world();
  return 0;
}
