#pragma once
/* Domain decomposition and subdomain -> (rank, GPU) placement.

   Parity with reference include/stencil/partition.hpp:
     RankPartition  (:23-144)   flat split: prime factors of n, largest first, always cut the longest axis
     NodePartition  (:148-310)  two-level split (nodes, then GPUs/node); cut the axis with the smallest interface,
                                interface = face area * (radius(+axis) + radius(-axis))
     Placement/Trivial/NodeAware (:312-864)
   Remainder rule (both): with size = q*dim + rem along an axis, the first `rem` subdomains get ceil, the rest floor.

   Differences by design:
   * no MPI: set-up collectives go through stencil::Comm (comm.hpp);
   * NodeAware takes an injectable bandwidth(cudaA, cudaB) function so placement is unit-testable on synthetic
     topologies without GPUs/NVML (BASELINE config 1);
   * the global index of node-local subdomain `id` on node `node` is sysIdx*nodeDim + node_idx(id) everywhere
     (the reference decodes with the global grid, which is only right for one node — SURVEY 2.8-2);
   * the comm matrix accumulates when two directions reach the same neighbour. */
#include <cassert>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <map>
#include <numeric>
#include <vector>

#include "stencil/comm.hpp"
#include "stencil/dim3.hpp"
#include "stencil/halo_geometry.hpp"
#include "stencil/logging.hpp"
#include "stencil/mat2d.hpp"
#include "stencil/mpi_topology.hpp"
#include "stencil/qap.hpp"
#include "stencil/radius.hpp"

namespace stencil {
namespace detail {

/* prime factors, largest first */
inline std::vector<int64_t> prime_factors_desc(int64_t n) {
  std::vector<int64_t> f;
  for (int64_t p = 2; p * p <= n; p += (p == 2 ? 1 : 2))
    while (n % p == 0) {
      f.push_back(p);
      n /= p;
    }
  if (n > 1) f.push_back(n);
  std::sort(f.begin(), f.end(), [](int64_t a, int64_t b) { return a > b; });
  return f;
}

inline int64_t div_ceil(int64_t n, int64_t d) { return (n + d - 1) / d; }

/* How the partitioners treat the x axis. The reference takes x first whenever its criterion ties (every cubic domain;
   partition.hpp:52-58, :183-189 there) and weighs all axes alike. x is the contiguous axis: cutting it creates faces
   made of 4-24 byte row fragments — one DRAM sector and one NVLink packet per row on B200, measured ~3x the cost per
   byte of a y or z face (profiles/README.md section 3) — and shortens the rows a 128-cell tile has to fill. Default
   here (TieBreak::Z): ties go z before y before x, and a cut through x counts kXCutPenalty times its size, so
   1024^3 on 8 GPUs becomes 1x2x4 boxes of 1024x512x256 (same halo volume as 2x2x2 cubes, no strided face on a link).
   TieBreak::X restores the reference's subdomain grids (also: env STENCIL_PARTITION_TIES=x). */
enum class TieBreak { Z, X };
constexpr int64_t kXCutPenalty = 3;
inline TieBreak &tie_break_storage() {
  static TieBreak t = [] {
    const char *e = std::getenv("STENCIL_PARTITION_TIES");
    return (e && (e[0] == 'x' || e[0] == 'X')) ? TieBreak::X : TieBreak::Z;
  }();
  return t;
}
/* wantLargest: the values are axis lengths and the longest is cut (RankPartition); else they are interface sizes and
   the smallest is cut (NodePartition) */
inline int pick_axis(int64_t x, int64_t y, int64_t z, bool wantLargest) {
  const bool b200 = tie_break_storage() == TieBreak::Z;
  /* compare x/penalty against y, z without division: scale the other two instead */
  const int64_t px = b200 ? kXCutPenalty : 1;
  const int64_t v[3] = {wantLargest ? x : -x * px, wantLargest ? y * px : -y, wantLargest ? z * px : -z};
  const int order[2][3] = {{2, 1, 0}, {0, 1, 2}};
  const int *o = order[tie_break_storage() == TieBreak::X ? 1 : 0];
  int best = o[0];
  for (int k = 1; k < 3; ++k)
    if (v[o[k]] > v[best]) best = o[k];
  return best;
}

inline int64_t linearize(const Dim3 &idx, const Dim3 &dim) {
  assert(idx.all_ge(0) && idx.all_lt(dim));
  return idx.x + dim.x * (idx.y + dim.y * idx.z);
}

inline Dim3 dimensionize(int64_t i, const Dim3 &dim) {
  assert(i >= 0 && i < dim.flatten());
  return Dim3(i % dim.x, (i / dim.x) % dim.y, i / (dim.x * dim.y));
}

/* sizes/origins under the "first rem get ceil" rule; ceilSz is the ceil size, rem = total % dim */
inline Dim3 part_size(const Dim3 &ceilSz, const Dim3 &rem, const Dim3 &idx) {
  Dim3 s = ceilSz;
  for (int a = 0; a < 3; ++a)
    if (rem[a] != 0 && idx[a] >= rem[a]) s[a] -= 1;
  return s;
}
inline Dim3 part_origin(const Dim3 &ceilSz, const Dim3 &rem, const Dim3 &idx) {
  Dim3 o = ceilSz * idx;
  for (int a = 0; a < 3; ++a)
    if (rem[a] != 0 && idx[a] >= rem[a]) o[a] -= idx[a] - rem[a];
  return o;
}

} // namespace detail
} // namespace stencil

/* tie-breaking order of both partitioners (process-wide; set before realize()) */
inline void set_partition_tie_break(stencil::detail::TieBreak t) { stencil::detail::tie_break_storage() = t; }
inline stencil::detail::TieBreak partition_tie_break() { return stencil::detail::tie_break_storage(); }

class RankPartition {
  Dim3 dim_;  // subdomains per axis
  Dim3 size_; // ceil subdomain size
  Dim3 rem_;  // input size % dim_

public:
  RankPartition() : dim_(1, 1, 1) {}
  RankPartition(const Dim3 &size, int64_t n) : dim_(1, 1, 1), size_(size) {
    for (int64_t f : stencil::detail::prime_factors_desc(n)) {
      const int a = stencil::detail::pick_axis(size_.x, size_.y, size_.z, /*wantLargest=*/true);
      size_[a] = stencil::detail::div_ceil(size_[a], f);
      dim_[a] *= f;
    }
    rem_ = size % dim_;
  }
  Dim3 dim() const { return dim_; }
  Dim3 subdomain_size(const Dim3 &idx) const { return stencil::detail::part_size(size_, rem_, idx); }
  Dim3 subdomain_origin(const Dim3 &idx) const { return stencil::detail::part_origin(size_, rem_, idx); }
  size_t linearize(const Dim3 &idx) const { return size_t(stencil::detail::linearize(idx, dim_)); }
  Dim3 dimensionize(int64_t i) const { return stencil::detail::dimensionize(i, dim_); }
};

class NodePartition {
  Dim3 sysDim_, nodeDim_;
  Dim3 size_, rem_;

  void split(Dim3 &dim, const Radius &radius, int64_t n) {
    for (int64_t f : stencil::detail::prime_factors_desc(n)) {
      const int64_t xi = size_.y * size_.z * int64_t(radius.x(1) + radius.x(-1));
      const int64_t yi = size_.x * size_.z * int64_t(radius.y(1) + radius.y(-1));
      const int64_t zi = size_.x * size_.y * int64_t(radius.z(1) + radius.z(-1));
      const int a = stencil::detail::pick_axis(xi, yi, zi, /*wantLargest=*/false);
      size_[a] = stencil::detail::div_ceil(size_[a], f);
      dim[a] *= f;
    }
  }

public:
  NodePartition() : sysDim_(1, 1, 1), nodeDim_(1, 1, 1) {}
  NodePartition(const Dim3 &size, const Radius &radius, int64_t nodes, int64_t gpus)
      : sysDim_(1, 1, 1), nodeDim_(1, 1, 1), size_(size) {
    split(sysDim_, radius, nodes);
    split(nodeDim_, radius, gpus);
    rem_ = size % (sysDim_ * nodeDim_);
  }
  Dim3 sys_dim() const noexcept { return sysDim_; }
  Dim3 node_dim() const noexcept { return nodeDim_; }
  Dim3 dim() const noexcept { return sysDim_ * nodeDim_; }
  Dim3 subdomain_size(const Dim3 &idx) const { return stencil::detail::part_size(size_, rem_, idx); }
  Dim3 subdomain_origin(const Dim3 &idx) const { return stencil::detail::part_origin(size_, rem_, idx); }
  Dim3 sys_idx(int64_t i) const { return stencil::detail::dimensionize(i, sysDim_); }
  Dim3 node_idx(int64_t i) const { return stencil::detail::dimensionize(i, nodeDim_); }
  Dim3 idx(int64_t i) const { return stencil::detail::dimensionize(i, dim()); }
};

enum class PlacementStrategy { NodeAware, Trivial };

class Placement {
public:
  virtual ~Placement() {}
  /* index (in the subdomain grid) of subdomain `i` of `rank` */
  virtual Dim3 get_idx(int rank, int i) = 0;
  virtual int get_rank(const Dim3 &idx) = 0;
  /* position of the subdomain inside its rank's domain list */
  virtual int get_subdomain_id(const Dim3 &idx) = 0;
  virtual int get_cuda(const Dim3 &idx) = 0;
  virtual Dim3 subdomain_size(const Dim3 &idx) = 0;
  virtual Dim3 subdomain_origin(const Dim3 &idx) = 0;
  virtual Dim3 dim() = 0;
  /* number of subdomains owned by `rank` */
  virtual int num_subdomains(int rank) = 0;
};

namespace stencil {
namespace detail {

/* shared storage: per linear subdomain index -> rank / id-in-rank / cuda device */
class PlacementTable : public Placement {
protected:
  Dim3 dim_;
  std::vector<int> rank_, id_, cuda_;
  std::vector<std::vector<Dim3>> idx_; // idx_[rank][id]

  void reset(const Dim3 &dim) {
    dim_ = dim;
    const size_t n = size_t(dim.flatten());
    rank_.assign(n, -1);
    id_.assign(n, -1);
    cuda_.assign(n, -1);
    idx_.clear();
  }
  void assign(const Dim3 &idx, int rank, int id, int cuda) {
    const size_t l = size_t(linearize(idx, dim_));
    rank_[l] = rank;
    id_[l] = id;
    cuda_[l] = cuda;
    if (idx_.size() <= size_t(rank)) idx_.resize(size_t(rank) + 1);
    if (idx_[size_t(rank)].size() <= size_t(id)) idx_[size_t(rank)].resize(size_t(id) + 1);
    idx_[size_t(rank)][size_t(id)] = idx;
  }

public:
  Dim3 get_idx(int rank, int i) override { return idx_.at(size_t(rank)).at(size_t(i)); }
  int get_rank(const Dim3 &idx) override { return rank_[size_t(linearize(idx, dim_))]; }
  int get_subdomain_id(const Dim3 &idx) override { return id_[size_t(linearize(idx, dim_))]; }
  int get_cuda(const Dim3 &idx) override { return cuda_[size_t(linearize(idx, dim_))]; }
  Dim3 dim() override { return dim_; }
  int num_subdomains(int rank) override { return size_t(rank) < idx_.size() ? int(idx_[size_t(rank)].size()) : 0; }
};

} // namespace detail
} // namespace stencil

/* Subdomain i (x-fastest linear order) goes to ranks in rank order, each rank taking as many consecutive
   subdomains as GPUs it contributes. */
class Trivial : public stencil::detail::PlacementTable {
  RankPartition part_;

public:
  Trivial(const Dim3 &size, MpiTopology &topo, const std::vector<int> &rankCudaIds) {
    stencil::Comm &comm = topo.comm();
    std::vector<int> counts(static_cast<size_t>(comm.size()));
    const int mine = int(rankCudaIds.size());
    comm.allgather(&mine, counts.data(), sizeof(int));
    std::vector<size_t> bytes(counts.size());
    for (size_t r = 0; r < counts.size(); ++r) bytes[r] = size_t(counts[r]) * sizeof(int);
    const int total = std::accumulate(counts.begin(), counts.end(), 0);
    std::vector<int> cudaOf(static_cast<size_t>(total));
    comm.allgatherv(rankCudaIds.data(), cudaOf.data(), bytes);
    build(size, counts, cudaOf);
  }
  /* communicator-free form: counts[r] GPUs per rank, cudaOf concatenated in rank order */
  Trivial(const Dim3 &size, const std::vector<int> &counts, const std::vector<int> &cudaOf) { build(size, counts, cudaOf); }

  Dim3 subdomain_size(const Dim3 &idx) override { return part_.subdomain_size(idx); }
  Dim3 subdomain_origin(const Dim3 &idx) override { return part_.subdomain_origin(idx); }

private:
  void build(const Dim3 &size, const std::vector<int> &counts, const std::vector<int> &cudaOf) {
    const int total = std::accumulate(counts.begin(), counts.end(), 0);
    part_ = RankPartition(size, total);
    reset(part_.dim());
    int i = 0;
    for (size_t r = 0; r < counts.size(); ++r)
      for (int id = 0; id < counts[r]; ++id, ++i) assign(part_.dimensionize(i), int(r), id, cudaOf[size_t(i)]);
  }
};

/* Node-aware placement: NodePartition, then per node a QAP between the node's subdomains (weights = halo elements
   exchanged, periodic wrap included) and the node's GPUs (distance = 1/bandwidth). */
class NodeAware : public stencil::detail::PlacementTable {
public:
  typedef std::function<double(int, int)> BandwidthFn;

private:
  NodePartition part_;
  std::vector<std::vector<size_t>> components_; // per node: QAP result (subdomain -> component)

public:
  /* bandwidth defaults to gpu_topo::bandwidth (declared in gpu_topology.hpp, resolved at link time) */
  NodeAware(const Dim3 &size, MpiTopology &topo, const Radius &radius, const std::vector<int> &rankCudaIds,
            BandwidthFn bandwidth = BandwidthFn());

  /* communicator-free form for tests and offline planning: `ranksOfNode[n]` lists the world ranks on node n, every
     rank contributes gpusPerRank devices, cudaOf[rank*gpusPerRank + g] is the device id. */
  NodeAware(const Dim3 &size, const Radius &radius, const std::vector<std::vector<int>> &ranksOfNode, int gpusPerRank,
            const std::vector<int> &cudaOf, BandwidthFn bandwidth) {
    std::vector<int> rankA, idA, cudaA;
    plan(size, radius, ranksOfNode, gpusPerRank, cudaOf, bandwidth, rankA, idA, cudaA);
    fill(ranksOfNode.size(), gpusPerRank * int(ranksOfNode.empty() ? 0 : ranksOfNode[0].size()), rankA, idA, cudaA);
  }

  Dim3 subdomain_size(const Dim3 &idx) override { return part_.subdomain_size(idx); }
  Dim3 subdomain_origin(const Dim3 &idx) override { return part_.subdomain_origin(idx); }
  const NodePartition &partition() const { return part_; }
  const std::vector<size_t> &components(size_t node) const { return components_[node]; }

  /* halo elements sent from a subdomain of size sz in direction dir */
  static double comm_cost(const Dim3 &dir, const Dim3 &sz, const Radius &radius) {
    return double(stencil::halo_extent(dir, sz, radius).flatten());
  }

  /* communication matrix between the node-local subdomains of node `sysIdx` */
  static Mat2D<double> comm_matrix(const NodePartition &part, const Dim3 &sysIdx, const Radius &radius) {
    const Dim3 nodeDim = part.node_dim();
    const Dim3 globalDim = part.dim();
    const int64_t n = nodeDim.flatten();
    Mat2D<double> comm(n, n, 0.0);
    for (int64_t i = 0; i < n; ++i) {
      const Dim3 src = sysIdx * nodeDim + part.node_idx(i);
      const Dim3 sz = part.subdomain_size(src);
      for (int z = -1; z <= 1; ++z)
        for (int y = -1; y <= 1; ++y)
          for (int x = -1; x <= 1; ++x) {
            const Dim3 dir(x, y, z);
            if (dir == Dim3(0, 0, 0) || radius.dir(dir * -1) == 0) continue;
            const Dim3 dst = (src + dir).wrapped(globalDim);
            const Dim3 rel = dst - sysIdx * nodeDim;
            if (!rel.all_ge(0) || !rel.all_lt(nodeDim)) continue; // neighbour lives on another node
            const int64_t j = stencil::detail::linearize(rel, nodeDim);
            if (j == i) continue;
            /* the message fills the receiver's halo on side -dir */
            comm.at(i, j) += comm_cost(dir * -1, sz, radius);
          }
    }
    return comm;
  }

private:
  void plan(const Dim3 &size, const Radius &radius, const std::vector<std::vector<int>> &ranksOfNode, int gpusPerRank,
            const std::vector<int> &cudaOf, const BandwidthFn &bandwidth, std::vector<int> &rankA, std::vector<int> &idA,
            std::vector<int> &cudaA) {
    const int numNodes = int(ranksOfNode.size());
    const int ranksPerNode = int(ranksOfNode[0].size());
    const int gpusPerNode = ranksPerNode * gpusPerRank;
    part_ = NodePartition(size, radius, numNodes, gpusPerNode);
    rankA.assign(size_t(numNodes * gpusPerNode), -1);
    idA = rankA;
    cudaA = rankA;
    components_.assign(size_t(numNodes), std::vector<size_t>());
    for (int node = 0; node < numNodes; ++node) {
      const std::vector<int> &ranks = ranksOfNode[size_t(node)];
      assert(int(ranks.size()) == ranksPerNode && "NodeAware needs the same number of ranks on every node");
      Mat2D<double> bw(gpusPerNode, gpusPerNode, 0.0);
      for (int ci = 0; ci < gpusPerNode; ++ci)
        for (int cj = 0; cj < gpusPerNode; ++cj) {
          const int di = cudaOf[size_t(ranks[size_t(ci / gpusPerRank)] * gpusPerRank + ci % gpusPerRank)];
          const int dj = cudaOf[size_t(ranks[size_t(cj / gpusPerRank)] * gpusPerRank + cj % gpusPerRank)];
          bw.at(ci, cj) = bandwidth(di, dj);
        }
      const Mat2D<double> comm = comm_matrix(part_, part_.sys_idx(node), radius);
      const Mat2D<double> dist = make_reciprocal(bw);
      /* exact for node sizes where n! is cheap, local search beyond */
      std::vector<size_t> comp = gpusPerNode <= 9 ? qap::solve(comm, dist) : qap::solve_catch(comm, dist);
      for (int id = 0; id < gpusPerNode; ++id) {
        const size_t c = comp[size_t(id)];
        const int rank = ranks[c / size_t(gpusPerRank)];
        const int g = int(c % size_t(gpusPerRank));
        const size_t slot = size_t(node * gpusPerNode + id);
        rankA[slot] = rank;
        idA[slot] = g;
        cudaA[slot] = cudaOf[size_t(rank * gpusPerRank + g)];
      }
      components_[size_t(node)] = comp;
    }
  }

  void fill(size_t numNodes, int gpusPerNode, const std::vector<int> &rankA, const std::vector<int> &idA,
            const std::vector<int> &cudaA) {
    reset(part_.dim());
    const Dim3 nodeDim = part_.node_dim();
    for (size_t node = 0; node < numNodes; ++node)
      for (int id = 0; id < gpusPerNode; ++id) {
        const size_t slot = node * size_t(gpusPerNode) + size_t(id);
        const Dim3 idx = part_.sys_idx(int64_t(node)) * nodeDim + part_.node_idx(id);
        assign(idx, rankA[slot], idA[slot], cudaA[slot]);
      }
  }
};
