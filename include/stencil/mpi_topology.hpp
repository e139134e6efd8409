#pragma once
/* MpiTopology: which ranks share this node.
   Parity: reference include/stencil/mpi_topology.hpp:7-96 (MPI_Comm_split_type(SHARED) + allgather). Here the
   node grouping comes from an allgather of hostnames over stencil::Comm; ranks of a node are listed in world-rank
   order and colocated_rank() is the position in that list. */
#include <string>
#include <vector>

#include "stencil/comm.hpp"

class MpiTopology {
  stencil::Comm *comm_;
  std::vector<std::string> hosts_;  // hostname of every rank
  std::vector<int> colocated_;      // world ranks on my node, ascending
  std::vector<int> nodeOfRank_;     // node number (order of first appearance) per rank
  int numNodes_;

public:
  explicit MpiTopology(stencil::Comm &comm);
  MpiTopology() : MpiTopology(stencil::world()) {}
  /* synthetic topology for tests: hosts[r] is the node name of rank r, `me` is this rank */
  MpiTopology(const std::vector<std::string> &hosts, int me);

  int rank() const noexcept { return me_; }
  int size() const noexcept { return int(hosts_.size()); }
  stencil::Comm &comm() const noexcept { return *comm_; }
  int colocated_rank() const noexcept;
  int colocated_size() const noexcept { return int(colocated_.size()); }
  const std::vector<int> &colocated_ranks() const noexcept { return colocated_; }
  bool colocated(int rank) const noexcept;
  int num_nodes() const noexcept { return numNodes_; }
  int node_of(int rank) const noexcept { return nodeOfRank_[rank]; }
  const std::string &host_of(int rank) const noexcept { return hosts_[rank]; }

private:
  int me_;
  void build();
};
