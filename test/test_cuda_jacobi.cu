/* Jacobi3D engine against the host loop, bit for bit, in all three schedules (fused push, reference-style
   overlap, no overlap), with 1..8 subdomains on GPU 0, uneven sizes, and on every visible GPU. */
#include "harness.hpp"
#include <cmath>
#include <cstring>
#include <vector>
#include "stencil/jacobi.hpp"

namespace {

std::vector<float> gather(stencil::Jacobi3D &j) {
  DistributedDomain &dd = j.domain();
  const Dim3 g = dd.size();
  std::vector<float> out(size_t(g.flatten()), -7.f);
  j.sync();
  for (LocalDomain &ld : dd.domains()) {
    const std::vector<unsigned char> b = ld.interior_to_host(0);
    const float *v = reinterpret_cast<const float *>(b.data());
    const Dim3 sz = ld.size(), o = ld.origin();
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x) out[size_t(((o.z + z) * g.y + o.y + y) * g.x + o.x + x)] = v[i++];
  }
  return out;
}

size_t mismatches(const Dim3 &size, stencil::Jacobi3D::Mode mode, const std::vector<int> &gpus, int iters, int tileZ = 32,
                  MethodFlags methods = MethodFlags::All, bool narrowTiles = true, bool flaggedFirst = true) {
  stencil::Jacobi3D::Options opt;
  opt.narrowTiles = narrowTiles;
  opt.flaggedFirst = flaggedFirst;
  opt.mode = mode;
  opt.gpus = gpus;
  opt.tileZ = tileZ;
  opt.methods = methods;
  stencil::Jacobi3D j(size_t(size.x), size_t(size.y), size_t(size.z), opt);
  std::vector<float> ref(size_t(size.flatten()), 0.5f), tmp;
  size_t bad = 0;
  for (int it = 0; it < iters; ++it) {
    j.step();
    stencil::jacobi_host_step(tmp, ref, size);
    ref.swap(tmp);
    if (it == 0 || it == iters - 1) {
      const std::vector<float> got = gather(j);
      for (size_t i = 0; i < got.size(); ++i) bad += std::memcmp(&got[i], &ref[i], 4) != 0;
    }
  }
  return bad;
}

int device_count() {
  int n = 0;
  cudaGetDeviceCount(&n);
  return n;
}

} // namespace

TEST_CASE("jacobi: the 3-instruction divide by six is correctly rounded", "[cuda][jacobi]") {
  std::vector<float> v;
  unsigned seed = 12345u;
  for (int i = 0; i < (1 << 20); ++i) {
    seed = seed * 1664525u + 1013904223u;
    /* sums of six values in [0,1]: cover [0,6] densely, plus a sweep of exponents */
    v.push_back(float(seed >> 8) * (6.0f / 16777216.0f));
  }
  for (int e = -100; e <= 100; ++e)
    for (int m = 0; m < 64; ++m) v.push_back(std::ldexp(1.0f + float(m) / 64.0f + 1e-7f * float(m), e));
  for (int i = 0; i < (1 << 18); ++i) {
    seed = seed * 1664525u + 1013904223u;
    unsigned bits = (seed & 0x007fffffu) | (unsigned(64 + (seed >> 26)) << 23); // exponents 2^-63 .. 2^0
    float f;
    std::memcpy(&f, &bits, 4);
    v.push_back(f);
  }
  REQUIRE(stencil::jacobi_div6_mismatches(v) == 0);
}

TEST_CASE("jacobi: both kernel variants agree with the host", "[cuda][jacobi]") {
  using M = stencil::Jacobi3D::Mode;
  for (bool tma : {true, false})
    for (M m : {M::Fused, M::NoOverlap}) {
      stencil::Jacobi3D::Options opt;
      opt.mode = m;
      opt.gpus = {0, 0};
      opt.tileZ = 8;
      opt.useTma = tma;
      const Dim3 size(150, 37, 21);
      stencil::Jacobi3D j(size_t(size.x), size_t(size.y), size_t(size.z), opt);
      std::vector<float> ref(size_t(size.flatten()), 0.5f), tmp;
      for (int it = 0; it < 5; ++it) {
        j.step();
        stencil::jacobi_host_step(tmp, ref, size);
        ref.swap(tmp);
      }
      const std::vector<float> got = gather(j);
      size_t bad = 0;
      for (size_t i = 0; i < got.size(); ++i) bad += std::memcmp(&got[i], &ref[i], 4) != 0;
      REQUIRE(bad == 0);
    }
}

TEST_CASE("jacobi: one subdomain, all schedules, bit exact", "[cuda][jacobi]") {
  using M = stencil::Jacobi3D::Mode;
  for (M m : {M::Fused, M::Overlap, M::NoOverlap}) {
    REQUIRE(mismatches(Dim3(40, 24, 20), m, {0}, 6) == 0);
    REQUIRE(mismatches(Dim3(130, 17, 9), m, {0}, 4, 4) == 0);   // partial x tile, partial row block, short z tiles
    REQUIRE(mismatches(Dim3(257, 33, 35), m, {0}, 3, 16) == 0); // three x tiles, last one a single column
  }
}

TEST_CASE("jacobi: rows that are not a multiple of the 128-cell tile use 32 x 64 remainder tiles", "[cuda][jacobi][narrow]") {
  using M = stencil::Jacobi3D::Mode;
  for (M m : {M::Fused, M::Overlap, M::NoOverlap}) {
    REQUIRE(mismatches(Dim3(133, 70, 20), m, {0}, 4, 8) == 0);    // one wide column + one narrow column, two narrow row blocks
    REQUIRE(mismatches(Dim3(70, 130, 12), m, {0}, 4, 4) == 0);    // narrow columns only (0, 32, 64), three row blocks
    REQUIRE(mismatches(Dim3(200, 65, 10), m, {0, 0}, 4, 8) == 0); // remainder 72: three narrow columns; y cut across subdomains
    REQUIRE(mismatches(Dim3(230, 20, 18), m, {0}, 3, 8) == 0);    // remainder 102 > 96: a second wide column instead
    REQUIRE(mismatches(Dim3(132, 64, 9), m, {0}, 3, 4) == 0);     // remainder is exactly one vector
  }
  /* same sizes with the remainder tiles switched off agree as well */
  REQUIRE(mismatches(Dim3(133, 70, 20), M::Fused, {0}, 4, 8, MethodFlags::All, false) == 0);
  /* x cut across subdomains (no wrap by load: x halos are pushed by narrow tiles too) */
  const stencil::detail::TieBreak saved = partition_tie_break();
  set_partition_tie_break(stencil::detail::TieBreak::X);
  const size_t bad = mismatches(Dim3(330, 70, 12), M::Fused, {0, 0}, 5, 4);
  set_partition_tie_break(saved);
  REQUIRE(bad == 0);
}

TEST_CASE("jacobi: thin layers at the top of a deep subdomain, automatic layer depth", "[cuda][jacobi][tail]") {
  using M = stencil::Jacobi3D::Mode;
  for (M m : {M::Fused, M::Overlap, M::NoOverlap}) {
    REQUIRE(mismatches(Dim3(140, 20, 70), m, {0}, 4, 8) == 0);       // 70 planes: 22 in three layers, 48 in six thin ones
    REQUIRE(mismatches(Dim3(140, 20, 140), m, {0, 0}, 4, 8) == 0);   // the same per subdomain, z faces pushed by a thin and a deep layer
    REQUIRE(mismatches(Dim3(133, 30, 81), m, {0}, 3, 16) == 0);      // odd plane count, narrow column
    REQUIRE(mismatches(Dim3(96, 40, 50), m, {0, 0}, 3, 0) == 0);     // depth chosen by the library
  }
}

TEST_CASE("jacobi: 2, 3, 4 and 8 subdomains on GPU 0", "[cuda][jacobi]") {
  using M = stencil::Jacobi3D::Mode;
  for (M m : {M::Fused, M::Overlap, M::NoOverlap}) {
    REQUIRE(mismatches(Dim3(40, 24, 20), m, {0, 0}, 6) == 0);
    REQUIRE(mismatches(Dim3(41, 23, 19), m, {0, 0, 0}, 5, 8) == 0);
    REQUIRE(mismatches(Dim3(64, 30, 22), m, {0, 0, 0, 0}, 5, 8) == 0);
    REQUIRE(mismatches(Dim3(36, 34, 30), m, {0, 0, 0, 0, 0, 0, 0, 0}, 5, 8) == 0);
  }
}

TEST_CASE("jacobi: legacy transports under the reference schedule", "[cuda][jacobi]") {
  using M = stencil::Jacobi3D::Mode;
  REQUIRE(mismatches(Dim3(40, 24, 20), M::Overlap, {0, 0}, 4, 32, MethodFlags::CudaMpi) == 0);
  REQUIRE(mismatches(Dim3(40, 24, 20), M::Overlap, {0, 0}, 4, 32, MethodFlags::CudaMemcpyPeer) == 0);
  REQUIRE(mismatches(Dim3(40, 24, 20), M::NoOverlap, {0, 0, 0, 0}, 4, 32, MethodFlags::CudaKernel | MethodFlags::CudaMpi) == 0);
}

TEST_CASE("jacobi: hot and cold spheres appear and the field stays in [0,1]", "[cuda][jacobi]") {
  stencil::Jacobi3D::Options opt;
  opt.gpus = {0};
  stencil::Jacobi3D j(60, 30, 30, opt);
  for (int i = 0; i < 10; ++i) j.step();
  float hot = -1, cold = -1, far = -1;
  REQUIRE(j.probe(Dim3(20, 15, 15), &hot));
  REQUIRE(j.probe(Dim3(40, 15, 15), &cold));
  REQUIRE(j.probe(Dim3(0, 0, 0), &far));
  REQUIRE(hot == 1.0f);
  REQUIRE(cold == 0.0f);
  REQUIRE(far == 0.5f); // ten steps cannot reach the corner
  const double sum = j.checksum();
  REQUIRE(sum > 0.45 * 60 * 30 * 30);
  REQUIRE(sum < 0.55 * 60 * 30 * 30);
  REQUIRE(j.launches() >= 10);
}

TEST_CASE("jacobi: random initial field and fingerprint are independent of partition and schedule", "[cuda][jacobi][random]") {
  using M = stencil::Jacobi3D::Mode;
  const Dim3 size(133, 40, 30);
  std::vector<float> ref(size_t(size.flatten())), tmp;
  for (int64_t z = 0; z < size.z; ++z)
    for (int64_t y = 0; y < size.y; ++y)
      for (int64_t x = 0; x < size.x; ++x) ref[size_t((z * size.y + y) * size.x + x)] = stencil::jacobi_random_value(7, Dim3(x, y, z), size);
  const int iters = 6;
  for (int it = 0; it < iters; ++it) {
    stencil::jacobi_host_step(tmp, ref, size);
    ref.swap(tmp);
  }
  std::vector<unsigned long long> prints;
  for (M m : {M::Fused, M::NoOverlap, M::Overlap})
    for (int nsub : {1, 2, 4}) {
      stencil::Jacobi3D::Options opt;
      opt.mode = m;
      opt.gpus.assign(size_t(nsub), 0);
      opt.tileZ = 8;
      stencil::Jacobi3D j(size_t(size.x), size_t(size.y), size_t(size.z), opt);
      j.step(); // something else first: init_random() must also work on an engine that has already run
      j.init_random(7);
      for (int it = 0; it < iters; ++it) j.step();
      const std::vector<float> got = gather(j);
      size_t bad = 0;
      for (size_t i = 0; i < got.size(); ++i) bad += std::memcmp(&got[i], &ref[i], 4) != 0;
      REQUIRE(bad == 0);
      prints.push_back(j.fingerprint());
    }
  for (unsigned long long p : prints) REQUIRE(p == prints[0]);
  /* ... and it does tell fields apart: one more step changes it */
  stencil::Jacobi3D::Options opt;
  opt.gpus = {0};
  stencil::Jacobi3D j(size_t(size.x), size_t(size.y), size_t(size.z), opt);
  j.init_random(7);
  for (int it = 0; it < iters + 1; ++it) j.step();
  REQUIRE(j.fingerprint() != prints[0]);
}

TEST_CASE("jacobi: every visible GPU, fused NVLink pushes", "[cuda][jacobi][multigpu]") {
  const int n = device_count();
  if (n < 2) {
    std::printf("  (1 GPU visible: multi-GPU fused path not exercised)\n");
    return;
  }
  std::vector<int> gpus;
  for (int i = 0; i < n; ++i) gpus.push_back(i);
  using M = stencil::Jacobi3D::Mode;
  {
    /* x cut across GPUs (the reference's partition): per-row sector stores over NVLink */
    const stencil::detail::TieBreak saved = partition_tie_break();
    set_partition_tie_break(stencil::detail::TieBreak::X);
    const size_t bad = mismatches(Dim3(70, 40, 36), M::Fused, gpus, 8, 8);
    /* ... and with wide tiles only, in natural tile order */
    const size_t badWide = mismatches(Dim3(70, 40, 36), M::Fused, gpus, 8, 8, MethodFlags::All, false, false);
    set_partition_tie_break(saved);
    REQUIRE(bad == 0);
    REQUIRE(badWide == 0);
  }
  REQUIRE(mismatches(Dim3(70, 40, 36), M::Fused, gpus, 8, 8) == 0);
  REQUIRE(mismatches(Dim3(333, 150, 90), M::Fused, gpus, 12, 8) == 0); // wide + narrow tile columns, several row blocks
  REQUIRE(mismatches(Dim3(70, 40, 36), M::Overlap, gpus, 4, 8) == 0);
  std::vector<int> twice = gpus;
  twice.insert(twice.end(), gpus.begin(), gpus.end());
  REQUIRE(mismatches(Dim3(70, 40, 36), M::Fused, twice, 8, 8) == 0);
}

TEST_CASE("jacobi: step_io uploads per-step parameters and reads back per-step results without stalling the loop", "[cuda][jacobi][io]") {
  const Dim3 size(60, 24, 20);
  const int iters = 7;
  for (int nsub : {1, 2}) {
    stencil::Jacobi3D::Options opt;
    opt.gpus.assign(size_t(nsub), 0);
    opt.tileZ = 8;
    stencil::Jacobi3D j(size_t(size.x), size_t(size.y), size_t(size.z), opt);
    float *params = nullptr, *results = nullptr;
    CUDA_RUNTIME(cudaHostAlloc(&params, iters * 2 * sizeof(float), cudaHostAllocDefault));
    CUDA_RUNTIME(cudaHostAlloc(&results, size_t(iters) * size_t(nsub) * sizeof(float), cudaHostAllocDefault));
    std::vector<float> ref(size_t(size.flatten()), 0.5f), tmp;
    std::vector<std::vector<float>> refs;
    const int64_t R = size.x / 10, hx = size.x / 3, cx = size.x * 2 / 3, cy = size.y / 2, cz = size.z / 2;
    for (int k = 0; k < iters; ++k) {
      /* each step has its own slot in pinned memory, so the host may run ahead of the copies */
      params[2 * k] = 1.0f + 0.25f * float(k);
      params[2 * k + 1] = -0.5f * float(k);
      j.step_io(params + 2 * k, 2 * sizeof(float), results + size_t(k) * size_t(nsub));
      stencil::jacobi_host_step(tmp, ref, size);
      for (int64_t z = 0; z < size.z; ++z)
        for (int64_t y = 0; y < size.y; ++y)
          for (int64_t x = 0; x < size.x; ++x) {
            const int64_t base = (y - cy) * (y - cy) + (z - cz) * (z - cz);
            float &v = tmp[size_t((z * size.y + y) * size.x + x)];
            if (base + (x - hx) * (x - hx) < (R + 1) * (R + 1))
              v = params[2 * k];
            else if (base + (x - cx) * (x - cx) < (R + 1) * (R + 1))
              v = params[2 * k + 1];
          }
      ref.swap(tmp);
      refs.push_back(ref);
    }
    const std::vector<float> got = gather(j); // synchronises, read-back stream included
    size_t bad = 0;
    for (size_t i = 0; i < got.size(); ++i) bad += std::memcmp(&got[i], &ref[i], 4) != 0;
    REQUIRE(bad == 0);
    size_t badResults = 0;
    for (int k = 0; k < iters; ++k)
      for (size_t di = 0; di < j.domain().domains().size(); ++di) {
        const Rect3 r = j.domain().domains()[di].get_compute_region();
        const Dim3 c((r.lo.x + r.hi.x) / 2, (r.lo.y + r.hi.y) / 2, (r.lo.z + r.hi.z) / 2);
        const float want = refs[size_t(k)][size_t((c.z * size.y + c.y) * size.x + c.x)];
        badResults += std::memcmp(&want, &results[size_t(k) * size_t(nsub) + di], 4) != 0;
      }
    REQUIRE(badResults == 0);
    cudaFreeHost(params);
    cudaFreeHost(results);
  }
}
