/* NodeAware placement over a live process group. Reference: include/stencil/partition.hpp:631-863. */
#include "stencil/partition.hpp"

#include "stencil/gpu_topology.hpp"

NodeAware::NodeAware(const Dim3 &size, MpiTopology &topo, const Radius &radius, const std::vector<int> &rankCudaIds,
                     BandwidthFn bandwidth) {
  stencil::Comm &comm = topo.comm();
  if (!bandwidth) bandwidth = [](int a, int b) { return gpu_topo::bandwidth(a, b); };

  const int gpusPerRank = int(rankCudaIds.size());
  const int numNodes = topo.num_nodes();
  const int ranksPerNode = topo.colocated_size();
  if (ranksPerNode * numNodes != comm.size()) LOG_FATAL("NodeAware placement needs the same number of ranks on every node");

  /* every rank contributes the same number of devices */
  std::vector<int> cudaOf(static_cast<size_t>(comm.size()) * size_t(gpusPerRank));
  comm.allgather(rankCudaIds.data(), cudaOf.data(), size_t(gpusPerRank) * sizeof(int));

  std::vector<std::vector<int>> ranksOfNode(static_cast<size_t>(numNodes), std::vector<int>());
  for (int r = 0; r < comm.size(); ++r) ranksOfNode[size_t(topo.node_of(r))].push_back(r);

  /* rank 0 decides (its view of the node's GPU fabric stands for every node, as in the reference), all follow */
  const int gpusPerNode = ranksPerNode * gpusPerRank;
  const size_t n = size_t(numNodes) * size_t(gpusPerNode);
  std::vector<int> rankA(n, -1), idA(n, -1), cudaA(n, -1);
  if (comm.rank() == 0) {
    plan(size, radius, ranksOfNode, gpusPerRank, cudaOf, bandwidth, rankA, idA, cudaA);
    LOG_INFO("NodeAware: " << part_.sys_dim() << "x" << part_.node_dim());
  } else {
    part_ = NodePartition(size, radius, numNodes, gpusPerNode);
  }
  comm.bcast(rankA.data(), n * sizeof(int), 0);
  comm.bcast(idA.data(), n * sizeof(int), 0);
  comm.bcast(cudaA.data(), n * sizeof(int), 0);
  fill(size_t(numNodes), gpusPerNode, rankA, idA, cudaA);
}
