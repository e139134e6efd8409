/* CopyPlan's chunk table, built on the host (no GPU needed): every row of every box is covered exactly once, chunks
   never cross a box, chunks that span planes cover whole planes, optional boxes come last so that leaving them out is
   a shorter table. The device side of the same plan is covered by test_cuda_pack / test_cuda_exchange. */
#include "harness.hpp"
#include <map>
#include <vector>
#include "stencil/copy_plan.hpp"

using stencil::BoxView;
using stencil::CopyChunk;
using stencil::CopyPlan;
using stencil::CopySeg;

TEST_CASE("copy plan chunks cover every row of every box exactly once", "[copyplan]") {
  /* fake addresses: the table is never dereferenced on the host */
  char *base = reinterpret_cast<char *>(0x10000000);
  const Dim3 arr(70, 50, 40);
  const BoxView a = BoxView::dense(base, arr, 4), b = BoxView::dense(base + (1 << 26), arr, 4);
  CopyPlan plan;
  const Dim3 boxes[][2] = {{Dim3(0, 0, 0), Dim3(64, 48, 3)},  // z face: long rows, many per plane
                           {Dim3(3, 0, 0), Dim3(3, 48, 38)},  // x face: thin rows
                           {Dim3(0, 2, 0), Dim3(64, 3, 38)},  // y face: few rows per plane -> chunks of whole planes
                           {Dim3(1, 1, 1), Dim3(3, 3, 3)},    // corner
                           {Dim3(0, 0, 0), Dim3(70, 50, 40)}}; // everything
  for (const auto &bx : boxes) plan.add(a, bx[0], b, bx[0], bx[1]);
  plan.add_optional(a, Dim3(5, 0, 0), b, Dim3(5, 0, 0), Dim3(2, 50, 40));
  plan.add(a, Dim3(9, 9, 9), b, Dim3(9, 9, 9), Dim3(0, 4, 4)); // empty: ignored
  REQUIRE(plan.num_segments() == 6);
  for (size_t chunkBytes : {size_t(32 * 1024), size_t(1024), size_t(1) << 20}) {
    plan.build_chunks(chunkBytes);
    const std::vector<CopySeg> &segs = plan.segments();
    const std::vector<CopyChunk> &chunks = plan.chunks();
    std::map<std::pair<unsigned, unsigned long long>, int> seen; // (segment, plane * ny + row) -> times covered
    size_t bad = 0;
    for (size_t c = 0; c < chunks.size(); ++c) {
      const CopyChunk &ch = chunks[c];
      if (ch.seg >= segs.size()) {
        ++bad;
        continue;
      }
      const CopySeg &s = segs[ch.seg];
      if (ch.nplanes > 1 && (ch.y0 != 0 || ch.nrows != s.ny)) ++bad;           // several planes: whole planes only
      if (ch.z + ch.nplanes > s.nz || ch.y0 + ch.nrows > s.ny || ch.nrows == 0) ++bad;
      if ((ch.seg == 5) != (c >= plan.num_required_chunks())) ++bad;            // the optional box (added sixth) is last
      for (unsigned p = 0; p < ch.nplanes; ++p)
        for (unsigned r = 0; r < ch.nrows; ++r) ++seen[{ch.seg, (unsigned long long)(ch.z + p) * s.ny + ch.y0 + r}];
    }
    size_t rows = 0;
    for (const CopySeg &s : segs) rows += size_t(s.ny) * s.nz;
    REQUIRE(bad == 0);
    REQUIRE(seen.size() == rows);
    size_t twice = 0;
    for (const auto &kv : seen) twice += kv.second != 1;
    REQUIRE(twice == 0);
    REQUIRE(plan.num_required_chunks() < chunks.size());
  }
  REQUIRE(plan.bytes() == size_t(4) * (64 * 48 * 3 + 3 * 48 * 38 + 64 * 3 * 38 + 27 + 70 * 50 * 40 + 2 * 50 * 40));
}
