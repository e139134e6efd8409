/* RankPartition / NodePartition. RankPartition expectations are those of the reference's
   test/test_cpu_partition.cpp; the NodePartition cases (untested there) use the values obtained by running the
   reference's code on the BASELINE configs. */
#include "harness.hpp"
#include "stencil/partition.hpp"

/* the expectations below are the reference's subdomain grids: ties cut x first */
struct ReferenceTies {
  stencil::detail::TieBreak saved = partition_tie_break();
  ReferenceTies() { set_partition_tie_break(stencil::detail::TieBreak::X); }
  ~ReferenceTies() { set_partition_tie_break(saved); }
};

TEST_CASE("rank partition even", "[partition]") {
  ReferenceTies ties;
  RankPartition p(Dim3(10, 5, 5), 2);
  REQUIRE(p.dim() == Dim3(2, 1, 1));
  REQUIRE(p.subdomain_size(Dim3(0, 0, 0)) == Dim3(5, 5, 5));
  REQUIRE(p.subdomain_size(Dim3(1, 0, 0)) == Dim3(5, 5, 5));
}

TEST_CASE("rank partition remainders", "[partition]") {
  ReferenceTies ties;
  RankPartition p(Dim3(10, 3, 1), 4);
  const int64_t sizes[4] = {3, 3, 2, 2}, origins[4] = {0, 3, 6, 8};
  for (int i = 0; i < 4; ++i) {
    REQUIRE(p.subdomain_size(Dim3(i, 0, 0)) == Dim3(sizes[i], 3, 1));
    REQUIRE(p.subdomain_origin(Dim3(i, 0, 0)) == Dim3(origins[i], 0, 0));
  }
  RankPartition q(Dim3(10, 5, 5), 3);
  REQUIRE(q.subdomain_size(Dim3(0, 0, 0)) == Dim3(4, 5, 5));
  REQUIRE(q.subdomain_size(Dim3(1, 0, 0)) == Dim3(3, 5, 5));
  REQUIRE(q.subdomain_size(Dim3(2, 0, 0)) == Dim3(3, 5, 5));
  RankPartition r(Dim3(13, 7, 7), 4);
  REQUIRE(r.subdomain_size(Dim3(0, 0, 0)) == Dim3(4, 7, 7));
  for (int i = 1; i < 4; ++i) REQUIRE(r.subdomain_size(Dim3(i, 0, 0)) == Dim3(3, 7, 7));
}

TEST_CASE("rank partition 3x3 origins", "[partition]") {
  ReferenceTies ties;
  RankPartition p(Dim3(10, 14, 2), 9);
  REQUIRE(p.dim() == Dim3(3, 3, 1));
  REQUIRE(p.subdomain_origin(Dim3(0, 0, 0)) == Dim3(0, 0, 0));
  REQUIRE(p.subdomain_origin(Dim3(1, 1, 0)) == Dim3(4, 5, 0));
  REQUIRE(p.subdomain_origin(Dim3(2, 2, 0)) == Dim3(7, 10, 0));
  /* tiles cover the domain exactly */
  int64_t cells = 0;
  for (int64_t i = 0; i < p.dim().flatten(); ++i) cells += p.subdomain_size(p.dimensionize(i)).flatten();
  REQUIRE(cells == 10 * 14 * 2);
  REQUIRE(p.linearize(p.dimensionize(7)) == 7);
}

static Radius faces(size_t r) {
  Radius x = Radius::constant(0);
  x.set_face(r);
  return x;
}

TEST_CASE("node partition weak-scaling shapes", "[partition]") {
  ReferenceTies ties;
  /* jacobi3d weak scaling, face radius 1 */
  REQUIRE(NodePartition(Dim3(512, 512, 512), faces(1), 1, 1).dim() == Dim3(1, 1, 1));
  NodePartition p2(Dim3(645, 645, 645), faces(1), 1, 2);
  REQUIRE(p2.dim() == Dim3(2, 1, 1));
  REQUIRE(p2.subdomain_size(Dim3(0, 0, 0)) == Dim3(323, 645, 645));
  REQUIRE(p2.subdomain_size(Dim3(1, 0, 0)) == Dim3(322, 645, 645));
  NodePartition p4(Dim3(813, 813, 813), faces(1), 1, 4);
  REQUIRE(p4.dim() == Dim3(2, 2, 1));
  REQUIRE(p4.subdomain_size(Dim3(0, 0, 0)) == Dim3(407, 407, 813));
  REQUIRE(p4.subdomain_size(Dim3(1, 1, 0)) == Dim3(406, 406, 813));
  NodePartition p8(Dim3(1024, 1024, 1024), faces(1), 1, 8);
  REQUIRE(p8.dim() == Dim3(2, 2, 2));
  REQUIRE(p8.subdomain_size(Dim3(1, 1, 1)) == Dim3(512, 512, 512));
  REQUIRE(p8.subdomain_origin(Dim3(1, 0, 1)) == Dim3(512, 0, 512));
}

TEST_CASE("node partition uneven radius changes the cut order", "[partition]") {
  ReferenceTies ties;
  Radius r = Radius::constant(1);
  r.dir(1, 0, 0) = 2; // +x=2, -x=1 => x interface weighted 3, y/z weighted 2
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 2).dim() == Dim3(1, 2, 1));
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 4).dim() == Dim3(1, 2, 2));
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 8).dim() == Dim3(2, 2, 2));
  REQUIRE(RankPartition(Dim3(1024, 1024, 1024), 2).dim() == Dim3(2, 1, 1));
}

TEST_CASE("node partition two levels and odd sizes", "[partition]") {
  ReferenceTies ties;
  NodePartition p(Dim3(1024, 1024, 1024), Radius::constant(3), 2, 4);
  REQUIRE(p.sys_dim() == Dim3(2, 1, 1));
  REQUIRE(p.node_dim() == Dim3(1, 2, 2));
  REQUIRE(p.dim() == Dim3(2, 2, 2));
  NodePartition q(Dim3(1023, 1024, 1025), Radius::constant(1), 1, 8);
  REQUIRE(q.subdomain_size(Dim3(0, 0, 0)) == Dim3(512, 512, 513));
  REQUIRE(q.subdomain_size(Dim3(1, 1, 1)) == Dim3(511, 512, 512));
  NodePartition s(Dim3(1000, 1000, 1000), Radius::constant(1), 1, 6);
  REQUIRE(s.dim() == Dim3(3, 2, 1));
  REQUIRE(s.subdomain_size(Dim3(0, 0, 0)) == Dim3(334, 500, 1000));
  REQUIRE(s.subdomain_size(Dim3(2, 1, 0)) == Dim3(333, 500, 1000));
  REQUIRE(s.sys_idx(0) == Dim3(0, 0, 0));
  REQUIRE(s.node_idx(4) == Dim3(1, 1, 0));
  REQUIRE(s.idx(5) == Dim3(2, 1, 0));
}

TEST_CASE("ties cut z before y before x by default, and the reference order on request", "[partition]") {
  REQUIRE(partition_tie_break() == stencil::detail::TieBreak::Z);
  const Radius r = Radius::constant(1);
  /* a cube on 2, 4, 8 GPUs: the contiguous axis is cut last */
  REQUIRE(NodePartition(Dim3(645, 645, 645), r, 1, 2).dim() == Dim3(1, 1, 2));
  REQUIRE(NodePartition(Dim3(813, 813, 813), r, 1, 4).dim() == Dim3(1, 2, 2));
  /* a cut through x counts three times its size: 8 GPUs get 1x2x4 boxes (the halo volume of 2x2x2 cubes, whole rows) */
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 8).dim() == Dim3(1, 2, 4));
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 8).subdomain_size(Dim3(0, 1, 3)) == Dim3(1024, 512, 256));
  REQUIRE(RankPartition(Dim3(1024, 1024, 1024), 8).dim() == Dim3(1, 2, 4));
  /* ... but x is still cut when it is the clearly better choice */
  REQUIRE(NodePartition(Dim3(4096, 256, 256), r, 1, 4).dim() == Dim3(4, 1, 1));
  REQUIRE(NodePartition(Dim3(1024, 1024, 1024), r, 1, 64).dim() == Dim3(2, 4, 8));
  REQUIRE(RankPartition(Dim3(64, 64, 64), 2).dim() == Dim3(1, 1, 2));
  REQUIRE(RankPartition(Dim3(64, 64, 64), 4).dim() == Dim3(1, 2, 2));
  /* no tie: the criterion decides, whatever the order */
  REQUIRE(RankPartition(Dim3(200, 64, 64), 2).dim() == Dim3(2, 1, 1));
  REQUIRE(RankPartition(Dim3(100, 64, 64), 2).dim() == Dim3(1, 1, 2)); // x is only 1.6x longer: not worth a strided face
  REQUIRE(NodePartition(Dim3(64, 512, 512), r, 1, 2).dim() == Dim3(1, 1, 2)); // y and z tie, x is the worst cut
  REQUIRE(NodePartition(Dim3(512, 512, 64), r, 1, 2).dim() == Dim3(1, 2, 1)); // x and y tie, z is the worst cut
  {
    ReferenceTies ties;
    REQUIRE(NodePartition(Dim3(645, 645, 645), r, 1, 2).dim() == Dim3(2, 1, 1));
    REQUIRE(RankPartition(Dim3(64, 64, 64), 4).dim() == Dim3(2, 2, 1));
  }
  REQUIRE(partition_tie_break() == stencil::detail::TieBreak::Z);
}

TEST_CASE("partitions tile the domain exactly for many sizes, counts and both tie policies", "[partition]") {
  const Radius r = Radius::constant(2);
  for (int policy = 0; policy < 2; ++policy) {
    const stencil::detail::TieBreak saved = partition_tie_break();
    set_partition_tie_break(policy ? stencil::detail::TieBreak::X : stencil::detail::TieBreak::Z);
    size_t bad = 0;
    const Dim3 sizes[] = {Dim3(17, 13, 11), Dim3(64, 64, 64), Dim3(645, 645, 645), Dim3(100, 7, 300), Dim3(5, 1024, 6), Dim3(813, 813, 813)};
    for (const Dim3 &size : sizes)
      for (int64_t n : {1, 2, 3, 4, 5, 6, 8, 12, 16, 30}) {
        for (int kind = 0; kind < 2; ++kind) {
          Dim3 dim;
          std::vector<Dim3> sz, org;
          auto collect = [&](auto &part) {
            dim = part.dim();
            for (int64_t z = 0; z < dim.z; ++z)
              for (int64_t y = 0; y < dim.y; ++y)
                for (int64_t x = 0; x < dim.x; ++x) {
                  sz.push_back(part.subdomain_size(Dim3(x, y, z)));
                  org.push_back(part.subdomain_origin(Dim3(x, y, z)));
                }
          };
          if (kind == 0) {
            RankPartition p(size, n);
            collect(p);
          } else {
            const int64_t nodes = n % 2 == 0 ? 2 : 1;
            NodePartition p(size, r, nodes, n / nodes);
            collect(p);
          }
          if (dim.flatten() != n) ++bad;
          int64_t cells = 0;
          for (size_t i = 0; i < sz.size(); ++i) {
            cells += sz[i].flatten();
            /* boxes of a tensor-product grid: sizes along an axis differ by at most one and origins are cumulative */
            for (int a = 0; a < 3; ++a) {
              if (sz[i][a] < size[a] / dim[a] || sz[i][a] > size[a] / dim[a] + 1) ++bad;
              if (org[i][a] < 0 || org[i][a] + sz[i][a] > size[a]) ++bad;
            }
            for (size_t j = i + 1; j < sz.size(); ++j) {
              bool overlap = true;
              for (int a = 0; a < 3; ++a) overlap = overlap && org[i][a] < org[j][a] + sz[j][a] && org[j][a] < org[i][a] + sz[i][a];
              if (overlap) ++bad;
            }
          }
          if (cells != size.flatten()) ++bad;
        }
      }
    set_partition_tie_break(saved);
    REQUIRE(bad == 0);
  }
}
