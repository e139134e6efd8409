/* Placement on synthetic 8-GPU topologies without GPUs, NVML or a process group (BASELINE config 1).
   The reference has no test for Trivial / NodeAware at all ("Gaps"). */
#include "harness.hpp"
#include <set>
#include "stencil/partition.hpp"

static std::vector<int> iota(int n) {
  std::vector<int> v(static_cast<size_t>(n));
  for (int i = 0; i < n; ++i) v[size_t(i)] = i;
  return v;
}

/* the expectations below are the reference's subdomain grids: ties cut x first */
struct ReferenceTies {
  stencil::detail::TieBreak saved = partition_tie_break();
  ReferenceTies() { set_partition_tie_break(stencil::detail::TieBreak::X); }
  ~ReferenceTies() { set_partition_tie_break(saved); }
};

TEST_CASE("trivial placement assigns in linear order", "[placement]") {
  ReferenceTies ties;
  /* two ranks: 3 GPUs and 1 GPU */
  Trivial t(Dim3(40, 10, 10), std::vector<int>{3, 1}, std::vector<int>{0, 1, 2, 0});
  REQUIRE(t.dim() == Dim3(4, 1, 1));
  REQUIRE(t.get_rank(Dim3(0, 0, 0)) == 0);
  REQUIRE(t.get_rank(Dim3(2, 0, 0)) == 0);
  REQUIRE(t.get_rank(Dim3(3, 0, 0)) == 1);
  REQUIRE(t.get_subdomain_id(Dim3(2, 0, 0)) == 2);
  REQUIRE(t.get_subdomain_id(Dim3(3, 0, 0)) == 0);
  REQUIRE(t.get_cuda(Dim3(1, 0, 0)) == 1);
  REQUIRE(t.get_idx(1, 0) == Dim3(3, 0, 0));
  REQUIRE(t.subdomain_size(Dim3(0, 0, 0)) == Dim3(10, 10, 10));
  REQUIRE(t.num_subdomains(0) == 3);
}

TEST_CASE("node-aware placement on a uniform NVSwitch box is the identity", "[placement]") {
  ReferenceTies ties;
  auto uniform = [](int a, int b) { return a == b ? 10.0 : 1.0; };
  NodeAware p(Dim3(512, 512, 512), Radius::constant(3), {{0}}, 8, iota(8), uniform);
  REQUIRE(p.dim() == Dim3(2, 2, 2));
  for (size_t i = 0; i < 8; ++i) REQUIRE(p.components(0)[i] == i);
  for (int i = 0; i < 8; ++i) {
    const Dim3 idx = p.partition().idx(i);
    REQUIRE(p.get_rank(idx) == 0);
    REQUIRE(p.get_subdomain_id(idx) == i);
    REQUIRE(p.get_cuda(idx) == i);
    REQUIRE(p.subdomain_size(idx) == Dim3(256, 256, 256));
  }
}

TEST_CASE("node-aware placement keeps heavy faces on fast links", "[placement]") {
  ReferenceTies ties;
  /* 4 GPUs in two fast pairs {0,1} and {2,3}. 400x300x8 with face radius 2 splits 2x2x1 (x first, then y) and the
     y faces (200x8) outweigh the x faces (150x8), so in linear order the heavy pairs are (0,2) and (1,3): the
     identity placement would put both on slow links and the solver has to move them. */
  auto islands = [](int a, int b) {
    if (a == b) return 100.0;
    return (a / 2 == b / 2) ? 50.0 : 1.0;
  };
  Radius r = Radius::constant(0);
  r.set_face(2);
  NodeAware p(Dim3(400, 300, 8), r, {{0}}, 4, iota(4), islands);
  REQUIRE(p.dim() == Dim3(2, 2, 1));
  const Dim3 dim = p.dim();
  REQUIRE(dim.flatten() == 4);
  /* find the axis pair with the largest shared face and check every such pair is on one island */
  const NodePartition &part = p.partition();
  const Mat2D<double> comm = NodeAware::comm_matrix(part, Dim3(0, 0, 0), r);
  double heavy = 0;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) heavy = std::max(heavy, comm.at(i, j));
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      if (comm.at(i, j) == heavy) {
        const int gi = p.get_cuda(part.node_idx(i)), gj = p.get_cuda(part.node_idx(j));
        REQUIRE(gi / 2 == gj / 2);
      }
  std::set<int> used;
  for (int i = 0; i < 4; ++i) used.insert(p.get_cuda(part.node_idx(i)));
  REQUIRE(used.size() == 4); // a bijection
}

TEST_CASE("node-aware placement across two nodes indexes the global grid", "[placement]") {
  ReferenceTies ties;
  auto uniform = [](int a, int b) { return a == b ? 10.0 : 1.0; };
  /* 2 nodes x 2 ranks x 2 GPUs; ranks 0,1 on node 0 and 2,3 on node 1 */
  std::vector<int> cudaOf = {0, 1, 2, 3, 0, 1, 2, 3};
  NodeAware p(Dim3(1024, 1024, 1024), Radius::constant(3), {{0, 1}, {2, 3}}, 2, cudaOf, uniform);
  REQUIRE(p.partition().sys_dim() == Dim3(2, 1, 1));
  REQUIRE(p.partition().node_dim() == Dim3(1, 2, 2));
  /* every subdomain with x index 0 belongs to node 0 (ranks 0/1), x index 1 to node 1 (ranks 2/3) */
  for (int z = 0; z < 2; ++z)
    for (int y = 0; y < 2; ++y) {
      REQUIRE(p.get_rank(Dim3(0, y, z)) <= 1);
      REQUIRE(p.get_rank(Dim3(1, y, z)) >= 2);
    }
  /* get_idx inverts (rank, id) */
  std::set<std::pair<int, int>> seen;
  for (int rank = 0; rank < 4; ++rank)
    for (int id = 0; id < 2; ++id) {
      const Dim3 idx = p.get_idx(rank, id);
      REQUIRE(p.get_rank(idx) == rank);
      REQUIRE(p.get_subdomain_id(idx) == id);
      REQUIRE(p.get_cuda(idx) == cudaOf[size_t(rank * 2 + id)]);
      seen.insert({rank, id});
    }
  REQUIRE(seen.size() == 8);
}

TEST_CASE("comm matrix counts both directions of a two-wide periodic axis", "[placement]") {
  NodePartition part(Dim3(64, 64, 64), Radius::constant(1), 1, 2);
  const Mat2D<double> comm = NodeAware::comm_matrix(part, Dim3(0, 0, 0), Radius::constant(1));
  /* +x and -x faces (32... the cut axis has 2 subdomains) plus the edges/corners that wrap to the same neighbour */
  const double face = 64.0 * 64.0;
  REQUIRE(comm.at(0, 1) >= 2 * face);
  REQUIRE(comm.at(0, 1) == comm.at(1, 0));
  REQUIRE(comm.at(0, 0) == 0);
}
