/* The tile tables the fused kernels walk (src/jacobi.cu make_tiles, src/astaroth.cu make_atiles), checked on the host:
   every cell of the region belongs to exactly one tile, the face / direction bits and the per-face tile counts the
   completion counters are compared against are right, tiles that push to another GPU come first and the cheap tiles
   last. A wrong count hangs a neighbour (flag never written) or releases it early, so this is the part of the
   hand-shake that can be wrong without any kernel being wrong. */
#include "harness.hpp"
#include <algorithm>
#include <cstdint>
#include <vector>
#include "stencil/astaroth.hpp"
#include "stencil/jacobi.hpp"

using stencil::detail::TileBox;

namespace {

/* cells of [lo,hi) covered by the boxes: every cell once? returns the number of cells covered != 1 times */
int64_t coverage_faults(const std::vector<TileBox> &t, const Dim3 &lo, const Dim3 &hi, int field) {
  const Dim3 e = hi - lo;
  std::vector<unsigned char> n(size_t(e.x * e.y * e.z), 0);
  int64_t outside = 0;
  for (const TileBox &b : t) {
    if (b.field != field) continue;
    for (int64_t z = b.lo.z; z < b.hi.z; ++z)
      for (int64_t y = b.lo.y; y < b.hi.y; ++y)
        for (int64_t x = b.lo.x; x < b.hi.x; ++x) {
          if (x < lo.x || x >= hi.x || y < lo.y || y >= hi.y || z < lo.z || z >= hi.z) {
            ++outside;
            continue;
          }
          ++n[size_t(((z - lo.z) * e.y + (y - lo.y)) * e.x + (x - lo.x))];
        }
  }
  int64_t bad = outside;
  for (unsigned char c : n) bad += c != 1;
  return bad;
}

} // namespace

TEST_CASE("jacobi tile table: coverage, face bits, counts, order", "[tiles][jacobi]") {
  const int tileZ = 16;
  struct Case {
    Dim3 sz;
    bool narrow;
  };
  const Case cases[] = {{Dim3(512, 64, 40), true},  {Dim3(645, 70, 33), true}, {Dim3(645, 70, 33), false}, {Dim3(813, 50, 17), true},
                        {Dim3(100, 33, 5), true},   {Dim3(129, 16, 16), true}, {Dim3(224, 65, 31), true},  {Dim3(225, 65, 31), true},
                        {Dim3(1024, 48, 20), true}, {Dim3(36, 130, 3), true}};
  for (const Case &c : cases) {
    const Dim3 lo(0, 0, 0), hi = c.sz;
    for (unsigned countMask : {0x3Fu, 0x30u, 0x0Cu, 0x00u, 0x3Cu}) {
      unsigned nFace[6];
      const std::vector<TileBox> t = stencil::detail::jacobi_tile_table(lo, hi, c.sz, tileZ, 0x3F, countMask, c.narrow, true, nFace);
      REQUIRE(!t.empty());
      REQUIRE(coverage_faults(t, lo, hi, 0) == 0);
      unsigned want[6] = {0, 0, 0, 0, 0, 0};
      bool narrowSeen = false;
      for (const TileBox &b : t) {
        /* a tile carries a face bit exactly when it owns cells of that face of the subdomain */
        const unsigned f = (b.lo.x == 0 ? 1u : 0u) | (b.hi.x == c.sz.x ? 2u : 0u) | (b.lo.y == 0 ? 4u : 0u) | (b.hi.y == c.sz.y ? 8u : 0u) |
                           (b.lo.z == 0 ? 16u : 0u) | (b.hi.z == c.sz.z ? 32u : 0u);
        REQUIRE(b.faces == f);
        for (int k = 0; k < 6; ++k)
          if ((f & countMask) >> k & 1u) ++want[k];
        narrowSeen |= b.shape == 1;
        if (b.shape == 1) REQUIRE(b.hi.x - b.lo.x <= 32);
        if (b.shape == 1) REQUIRE(c.narrow);
      }
      for (int k = 0; k < 6; ++k) REQUIRE(nFace[k] == want[k]);
      /* a remainder of at most 96 cells is covered by narrow tiles, anything else by wide ones */
      const int64_t rem = c.sz.x % 128;
      REQUIRE(narrowSeen == (c.narrow && rem != 0 && rem <= 96));
      /* order: z-face pushers, then the other pushers, then everything else; no pusher after a non-pusher */
      int stage = 0;
      for (const TileBox &b : t) {
        const int p = (b.faces & countMask & 0x30u) ? 0 : ((b.faces & countMask) ? 1 : 2);
        REQUIRE(p >= stage);
        stage = std::max(stage, p);
      }
      if (countMask) {
        /* cheap tiles (at most half the rows or half the planes of a full tile) that push nothing close the table */
        bool cheapSeen = false;
        for (const TileBox &b : t) {
          if (b.faces & countMask) continue;
          const int64_t rows = b.shape ? 64 : 16;
          const bool cheap = 2 * (b.hi.y - b.lo.y) <= rows || 2 * (b.hi.z - b.lo.z) <= tileZ;
          if (cheapSeen) REQUIRE(cheap);
          cheapSeen |= cheap;
        }
      }
    }
    /* natural order keeps (z, y, x) order and the same set of tiles */
    unsigned nA[6], nB[6];
    const auto a = stencil::detail::jacobi_tile_table(lo, hi, c.sz, tileZ, 0x3F, 0x3F, c.narrow, false, nA);
    const auto b = stencil::detail::jacobi_tile_table(lo, hi, c.sz, tileZ, 0x3F, 0x3F, c.narrow, true, nB);
    REQUIRE(a.size() == b.size());
    for (int k = 0; k < 6; ++k) REQUIRE(nA[k] == nB[k]);
    for (size_t i = 1; i < a.size(); ++i) REQUIRE(a[i].lo.z >= a[i - 1].lo.z);
  }
}

TEST_CASE("jacobi tile table: sub-regions (interior / exterior slabs)", "[tiles][jacobi]") {
  const Dim3 sz(645, 70, 33);
  const Dim3 los[] = {Dim3(1, 1, 1), Dim3(0, 0, 0), Dim3(644, 0, 0), Dim3(0, 69, 0), Dim3(130, 17, 16), Dim3(600, 3, 2)};
  const Dim3 his[] = {Dim3(644, 69, 32), Dim3(1, 70, 33), Dim3(645, 70, 33), Dim3(645, 70, 33), Dim3(300, 40, 33), Dim3(645, 66, 31)};
  for (size_t i = 0; i < sizeof(los) / sizeof(los[0]); ++i)
    for (bool narrow : {false, true}) {
      unsigned nFace[6];
      const auto t = stencil::detail::jacobi_tile_table(los[i], his[i], sz, 16, 0, 0, narrow, true, nFace);
      REQUIRE(coverage_faults(t, los[i], his[i], 0) == 0);
      for (const TileBox &b : t) {
        REQUIRE(b.faces == 0u); // non-fused launches push nothing
        REQUIRE(b.hi.x > b.lo.x);
        REQUIRE(b.hi.y > b.lo.y);
        REQUIRE(b.hi.z > b.lo.z);
      }
      for (int k = 0; k < 6; ++k) REQUIRE(nFace[k] == 0u);
    }
  /* an empty region has no tiles */
  unsigned nFace[6];
  REQUIRE(stencil::detail::jacobi_tile_table(Dim3(5, 5, 5), Dim3(5, 9, 9), sz, 16, 0x3F, 0x3F, true, true, nFace).empty());
}

TEST_CASE("astaroth tile table: coverage per field, direction bits, counts, order, balanced chunks", "[tiles][astaroth]") {
  struct Case {
    Dim3 sz;
    int fields, vecN, tileZ;
    bool narrow;
  };
  const Case cases[] = {{Dim3(256, 64, 64), 1, 4, 32, true},  {Dim3(162, 50, 162), 2, 2, 32, true}, {Dim3(162, 50, 162), 2, 2, 32, false},
                        {Dim3(100, 33, 7), 3, 4, 3, true},    {Dim3(645, 20, 35), 1, 4, 32, true},  {Dim3(80, 70, 40), 8, 2, 16, true},
                        {Dim3(48, 16, 9), 1, 2, 32, true},    {Dim3(130, 130, 33), 2, 4, 16, true}};
  for (const Case &c : cases) {
    const Dim3 lo(0, 0, 0), hi = c.sz;
    for (unsigned flagMask : {0xFFu, 0x00u, 0x0Cu, 0x03u, 0xF3u}) {
      unsigned nDir[8];
      int tz = 0;
      const auto t = stencil::detail::astaroth_tile_table(lo, hi, c.sz, c.fields, c.tileZ, c.vecN, 0xFF, flagMask, c.narrow, true, nDir, &tz);
      /* chunks: as many as the requested depth gives, of nearly the same depth, never shallower than the radius */
      const int64_t chunks = (c.sz.z + c.tileZ - 1) / c.tileZ;
      REQUIRE(tz >= 3);
      REQUIRE(tz <= std::max(c.tileZ, 3));
      REQUIRE(int64_t(tz) * chunks >= c.sz.z);
      REQUIRE(int64_t(tz - 1) * chunks < c.sz.z || tz == 3);
      for (int f = 0; f < c.fields; ++f) REQUIRE(coverage_faults(t, lo, hi, f) == 0);
      unsigned want[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (const TileBox &b : t) {
        REQUIRE(b.field >= 0);
        REQUIRE(b.field < c.fields);
        const bool ym = b.lo.y < 3, yp = b.hi.y > c.sz.y - 3, zm = b.lo.z < 3, zp = b.hi.z > c.sz.z - 3;
        const unsigned d = (ym ? 1u : 0u) | (yp ? 2u : 0u) | (zm ? 4u : 0u) | (zp ? 8u : 0u) | (ym && zm ? 16u : 0u) | (ym && zp ? 32u : 0u) |
                           (yp && zm ? 64u : 0u) | (yp && zp ? 128u : 0u);
        REQUIRE(b.faces == d);
        for (int k = 0; k < 8; ++k)
          if ((d & flagMask) >> k & 1u) ++want[k];
        if (b.shape == 1) REQUIRE(b.hi.x - b.lo.x <= 8 * c.vecN);
        if (b.shape == 1) REQUIRE(c.narrow);
      }
      for (int k = 0; k < 8; ++k) REQUIRE(nDir[k] == want[k]);
      int stage = 0;
      bool cheapSeen = false;
      for (const TileBox &b : t) {
        const int p = (b.faces & flagMask & 0x0Cu) ? 0 : ((b.faces & flagMask) ? 1 : 2);
        REQUIRE(p >= stage);
        stage = std::max(stage, p);
        if (p == 2) {
          const bool cheap = 2 * (b.hi.y - b.lo.y) <= (b.shape ? 64 : 16) || 2 * (b.hi.z - b.lo.z) <= tz;
          if (cheapSeen) REQUIRE(cheap);
          cheapSeen |= cheap;
        }
      }
    }
    /* without pushes (dirMask 0) no tile carries a direction and nothing is counted */
    unsigned nDir[8];
    for (const TileBox &b : stencil::detail::astaroth_tile_table(lo, hi, c.sz, c.fields, c.tileZ, c.vecN, 0, 0, c.narrow, true, nDir, nullptr))
      REQUIRE(b.faces == 0u);
    for (int k = 0; k < 8; ++k) REQUIRE(nDir[k] == 0u);
  }
}

TEST_CASE("astaroth tile table: sub-regions", "[tiles][astaroth]") {
  const Dim3 sz(162, 50, 162);
  const Dim3 los[] = {Dim3(3, 3, 3), Dim3(0, 0, 0), Dim3(159, 0, 0), Dim3(0, 47, 3), Dim3(70, 20, 100)};
  const Dim3 his[] = {Dim3(159, 47, 159), Dim3(3, 50, 162), Dim3(162, 50, 162), Dim3(162, 50, 159), Dim3(140, 40, 162)};
  for (size_t i = 0; i < sizeof(los) / sizeof(los[0]); ++i)
    for (int vecN : {2, 4})
      for (bool narrow : {false, true}) {
        const auto t = stencil::detail::astaroth_tile_table(los[i], his[i], sz, 2, 32, vecN, 0, 0, narrow, false, nullptr, nullptr);
        for (int f = 0; f < 2; ++f) REQUIRE(coverage_faults(t, los[i], his[i], f) == 0);
      }
}

TEST_CASE("jacobi tile table: thin layers at the top of the subdomain close the table", "[tiles][jacobi]") {
  const Dim3 sz(300, 40, 323), lo(0, 0, 0);
  for (unsigned countMask : {0x00u, 0x30u, 0x3Cu}) {
    unsigned nFace[6];
    const auto t = stencil::detail::jacobi_tile_table(lo, sz, sz, 32, 0x3F, countMask, true, true, nFace, 48, 8);
    REQUIRE(coverage_faults(t, lo, sz, 0) == 0);
    int64_t deepest = 0, thinPlanes = 0;
    for (const TileBox &b : t) {
      const int64_t d = b.hi.z - b.lo.z;
      deepest = std::max(deepest, d);
      REQUIRE(d <= 32);
      if (b.lo.z >= 323 - 48) REQUIRE(d <= 8);        // the top 48 planes: 8-plane layers
      if (b.lo.z < 323 - 48) REQUIRE(b.hi.z <= 323 - 48); // no layer straddles the boundary
      if (b.lo.z < 323 - 48) REQUIRE(d >= 18);        // the bulk (275 planes) in 9 even layers of 32 with one odd rest
      if (b.lo.x == 0 && b.lo.y == 0 && d <= 8) thinPlanes += d;
    }
    REQUIRE(deepest == 32);
    REQUIRE(thinPlanes == 48);
    /* the table ends with thin tiles (those that push to another GPU excepted: they go first) */
    const TileBox &last = t.back();
    REQUIRE(last.hi.z - last.lo.z <= 8);
    REQUIRE((last.faces & countMask) == 0u);
  }
  /* a region too shallow for bulk + tail keeps uniform layers */
  unsigned nFace[6];
  const Dim3 small(128, 16, 60);
  for (const TileBox &b : stencil::detail::jacobi_tile_table(lo, small, small, 32, 0, 0, true, true, nFace, 48, 8))
    REQUIRE((b.hi.z - b.lo.z == 32 || b.hi.z == 60));
}

TEST_CASE("astaroth tile table: thin chunks at the top of the subdomain close the table", "[tiles][astaroth]") {
  const Dim3 sz(130, 40, 256), lo(0, 0, 0);
  for (unsigned flagMask : {0x00u, 0x0Cu, 0xFFu}) {
    unsigned nDir[8];
    int tz = 0;
    const auto t = stencil::detail::astaroth_tile_table(lo, sz, sz, 3, 48, 2, 0xFF, flagMask, true, true, nDir, &tz, 16, 8);
    REQUIRE(tz == 48); // 240 planes of bulk in 5 chunks
    for (int f = 0; f < 3; ++f) REQUIRE(coverage_faults(t, lo, sz, f) == 0);
    for (const TileBox &b : t) {
      const int64_t d = b.hi.z - b.lo.z;
      if (b.lo.z >= 240) REQUIRE(d == 8);
      if (b.lo.z < 240) REQUIRE(d == 48);
    }
    REQUIRE(t.back().hi.z - t.back().lo.z == 8);
    REQUIRE((t.back().faces & flagMask) == 0u);
  }
  /* too shallow for bulk + tail: balanced chunks only */
  unsigned nDir[8];
  int tz = 0;
  const Dim3 small(64, 16, 100);
  for (const TileBox &b : stencil::detail::astaroth_tile_table(lo, small, small, 1, 48, 2, 0, 0, true, true, nDir, &tz, 16, 8)) REQUIRE(b.hi.z - b.lo.z >= 32);
  REQUIRE(tz == 34);
}
