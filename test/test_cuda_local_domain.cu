/* LocalDomain: geometry tables, padded layout, accessors, host transfers, swap, a 27-point stencil on 10^3.
   Ground covered by the reference's test_cuda_local_domain.cu (halo_pos/halo_extent tables for 30x40x50 r=4,
   x-leaning radius, 27-point stencil: corners 8, edges 12, faces 18, centre 27). */
#include "harness.hpp"
#include <cstring>
#include <vector>
#include "stencil/local_domain.cuh"

TEST_CASE("halo_pos / halo_extent tables, constant radius", "[cuda][local_domain]") {
  LocalDomain ld(Dim3(30, 40, 50), Dim3(0, 0, 0), 0);
  ld.set_radius(4);
  auto h = ld.add_data<float>();
  ld.realize();
  REQUIRE(ld.get_curr(h) != nullptr);
  REQUIRE(ld.get_next(h) != nullptr);
  REQUIRE(ld.get_curr(h) != ld.get_next(h));
  REQUIRE(ld.raw_size() == Dim3(38, 48, 58));
  const int64_t sz[3] = {30, 40, 50};
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; ++x) {
        const Dim3 dir(x, y, z);
        const Dim3 ext = ld.halo_extent(dir), ph = ld.halo_pos(dir, true), pi = ld.halo_pos(dir, false);
        for (int a = 0; a < 3; ++a) {
          REQUIRE(ext[a] == (dir[a] == 0 ? sz[a] : 4));
          REQUIRE(ph[a] == (dir[a] < 0 ? 0 : (dir[a] == 0 ? 4 : sz[a] + 4)));
          REQUIRE(pi[a] == (dir[a] < 0 ? 4 : (dir[a] == 0 ? 4 : sz[a])));
        }
        REQUIRE(ld.halo_bytes(dir, 0) == ext.flatten() * 4);
      }
  /* spot values spelled out */
  REQUIRE(ld.halo_pos(Dim3(1, 0, 0), true) == Dim3(34, 4, 4));
  REQUIRE(ld.halo_pos(Dim3(0, 1, 1), false) == Dim3(4, 40, 50));
  REQUIRE(ld.halo_extent(Dim3(1, 1, 0)) == Dim3(4, 4, 50));
  REQUIRE(ld.halo_extent(Dim3(-1, -1, -1)) == Dim3(4, 4, 4));
}

TEST_CASE("x-leaning radius geometry and coordinates", "[cuda][local_domain]") {
  LocalDomain ld(Dim3(30, 40, 50), Dim3(100, 200, 300), 0);
  Radius r = Radius::constant(0);
  r.dir(1, 0, 0) = 3;
  r.dir(-1, 0, 0) = 2;
  ld.set_radius(r);
  ld.add_data<double>();
  ld.realize();
  REQUIRE(ld.raw_size() == Dim3(35, 40, 50));
  REQUIRE(ld.halo_extent(Dim3(1, 0, 0)) == Dim3(3, 40, 50));
  REQUIRE(ld.halo_extent(Dim3(-1, 0, 0)) == Dim3(2, 40, 50));
  REQUIRE(ld.halo_extent(Dim3(0, 1, 0)) == Dim3(30, 0, 50));
  REQUIRE(ld.halo_pos(Dim3(1, 0, 0), true) == Dim3(32, 0, 0));
  REQUIRE(ld.halo_pos(Dim3(1, 0, 0), false) == Dim3(30, 0, 0));
  REQUIRE(ld.halo_pos(Dim3(-1, 0, 0), true) == Dim3(0, 0, 0));
  REQUIRE(ld.halo_pos(Dim3(-1, 0, 0), false) == Dim3(2, 0, 0));
  REQUIRE(ld.get_compute_region() == Rect3(Dim3(100, 200, 300), Dim3(130, 240, 350)));
  REQUIRE(ld.get_full_region() == Rect3(Dim3(98, 200, 300), Dim3(133, 240, 350)));
  REQUIRE(ld.halo_coords(Dim3(1, 0, 0), true) == Rect3(Dim3(130, 200, 300), Dim3(133, 240, 350)));
  REQUIRE(ld.halo_coords(Dim3(-1, 0, 0), true) == Rect3(Dim3(98, 200, 300), Dim3(100, 240, 350)));
  REQUIRE(ld.halo_coords(Dim3(-1, 0, 0), false) == Rect3(Dim3(100, 200, 300), Dim3(102, 240, 350)));
}

TEST_CASE("padded layout: aligned interior rows, dense host copies", "[cuda][local_domain]") {
  LocalDomain ld(Dim3(10, 3, 2), Dim3(0, 0, 0), 0);
  ld.set_radius(1);
  auto hf = ld.add_data<float>();
  auto hc = ld.add_data<char>();
  auto hd = ld.add_data<double>();
  ld.realize();
  for (size_t q = 0; q < 3; ++q) {
    const size_t es = ld.elem_size(q);
    /* interior x=0 of the first row: one radius past logical element 0 */
    const uintptr_t first = uintptr_t(ld.curr_data(q)) + es * 1;
    REQUIRE(first % 128 == 0);
    REQUIRE((size_t(ld.pitch(q).x) * es) % 128 == 0);
    REQUIRE(ld.pitch(q).x >= ld.raw_size().x);
    REQUIRE(ld.pitch(q).y == ld.raw_size().y);
  }
  (void)hc;
  (void)hd;
  /* host -> device -> host round trip through the padded rows is dense and exact */
  const Dim3 raw = ld.raw_size();
  std::vector<float> in(size_t(raw.flatten()));
  for (size_t i = 0; i < in.size(); ++i) in[i] = float(i) * 0.5f;
  ld.region_from_host(in.data(), Dim3(0, 0, 0), raw, 0);
  std::vector<unsigned char> all = ld.quantity_to_host(0);
  REQUIRE(all.size() == in.size() * sizeof(float));
  REQUIRE(std::memcmp(all.data(), in.data(), all.size()) == 0);
  std::vector<unsigned char> inner = ld.interior_to_host(0);
  REQUIRE(inner.size() == size_t(10 * 3 * 2) * sizeof(float));
  const float *fi = reinterpret_cast<const float *>(inner.data());
  REQUIRE(fi[0] == in[size_t((1 * raw.y + 1) * raw.x + 1)]);
  /* accessor addresses the same storage */
  Accessor<float> acc = ld.get_curr_accessor(hf);
  REQUIRE(acc.origin() == Dim3(-1, -1, -1));
  REQUIRE(acc.pitch() == ld.pitch(0));
  float v = -1;
  CUDA_RUNTIME(cudaMemcpy(&v, &acc[Dim3(0, 0, 0)], sizeof(float), cudaMemcpyDeviceToHost));
  REQUIRE(v == fi[0]);
  /* swap is pointer parity */
  float *c0 = ld.get_curr(hf), *n0 = ld.get_next(hf);
  void **t0 = ld.dev_curr_datas();
  ld.swap();
  REQUIRE(ld.get_curr(hf) == n0);
  REQUIRE(ld.get_next(hf) == c0);
  REQUIRE(ld.dev_curr_datas() != t0);
  void *tab[3];
  CUDA_RUNTIME(cudaMemcpy(tab, ld.dev_curr_datas(), sizeof(tab), cudaMemcpyDeviceToHost));
  REQUIRE(tab[0] == (void *)n0);
  size_t es[3];
  CUDA_RUNTIME(cudaMemcpy(es, ld.dev_elem_sizes(), sizeof(es), cudaMemcpyDeviceToHost));
  REQUIRE(es[0] == 4);
  REQUIRE(es[1] == 1);
  REQUIRE(es[2] == 8);
}

__global__ void fill_region(Accessor<float> dst, Rect3 reg, float v) {
  for (int64_t z = reg.lo.z + blockIdx.z; z < reg.hi.z; z += gridDim.z)
    for (int64_t y = reg.lo.y + threadIdx.y; y < reg.hi.y; y += blockDim.y)
      for (int64_t x = reg.lo.x + threadIdx.x; x < reg.hi.x; x += blockDim.x) dst.at(x, y, z) = v;
}

__global__ void sum27(Accessor<float> dst, Accessor<float> src, Rect3 reg) {
  for (int64_t z = reg.lo.z + blockIdx.z; z < reg.hi.z; z += gridDim.z)
    for (int64_t y = reg.lo.y + threadIdx.y; y < reg.hi.y; y += blockDim.y)
      for (int64_t x = reg.lo.x + threadIdx.x; x < reg.hi.x; x += blockDim.x) {
        float s = 0;
        for (int dz = -1; dz <= 1; ++dz)
          for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) s += src.at(x + dx, y + dy, z + dz);
        dst.at(x, y, z) = s;
      }
}

TEST_CASE("27-point stencil over a zero halo", "[cuda][local_domain]") {
  LocalDomain ld(Dim3(10, 10, 10), Dim3(0, 0, 0), 0);
  ld.set_radius(1);
  auto h = ld.add_data<float>();
  ld.realize();
  const dim3 grid(1, 1, 12), block(16, 16, 1);
  fill_region<<<grid, block>>>(ld.get_curr_accessor(h), ld.get_full_region(), 0.f);
  fill_region<<<grid, block>>>(ld.get_curr_accessor(h), ld.get_compute_region(), 1.f);
  sum27<<<grid, block>>>(ld.get_next_accessor(h), ld.get_curr_accessor(h), ld.get_compute_region());
  CUDA_RUNTIME(cudaDeviceSynchronize());
  ld.swap();
  const std::vector<unsigned char> bytes = ld.interior_to_host(0);
  const float *v = reinterpret_cast<const float *>(bytes.data());
  auto at = [&](int x, int y, int z) { return v[(z * 10 + y) * 10 + x]; };
  REQUIRE(at(0, 0, 0) == 8);
  REQUIRE(at(9, 9, 9) == 8);
  REQUIRE(at(5, 0, 0) == 12);
  REQUIRE(at(0, 9, 4) == 12);
  REQUIRE(at(5, 5, 0) == 18);
  REQUIRE(at(9, 3, 3) == 18);
  REQUIRE(at(5, 5, 5) == 27);
}
