/* The user-level loop the README of the reference teaches, end to end with user kernels and Accessors:
   init (ripple field, as the reference's test_derivative.cu:12-38) -> exchange -> a centred x/y/z difference
   written to `next` over get_interior() and get_exterior() -> swap -> read back. The ripple field
   f(p) = sum_a (p_a + r[p_a mod 4]), r = {0, .25, 0, -.25}, has the exact centred difference
   1 + (r[(p+1)%4] - r[(p-1)%4]) / 2 per axis away from the periodic seam, so every cell whose stencil crosses a
   subdomain boundary proves that the halo held the neighbour's values. */
#include "harness.hpp"
#include <cstring>
#include <vector>
#include "stencil/stencil.hpp"

namespace {

__host__ __device__ inline float ripple_at(int64_t v) {
  const int k = int(v & 3);
  return k == 1 ? 0.25f : (k == 3 ? -0.25f : 0.f);
}
__host__ __device__ inline float field(const Dim3 &p) {
  return float(p.x) + ripple_at(p.x) + float(p.y) + ripple_at(p.y) + float(p.z) + ripple_at(p.z);
}

__global__ void ripple_init(Accessor<float> dst, Rect3 box) {
  for (int64_t z = box.lo.z + blockIdx.z; z < box.hi.z; z += gridDim.z)
    for (int64_t y = box.lo.y + blockIdx.y * blockDim.y + threadIdx.y; y < box.hi.y; y += gridDim.y * blockDim.y)
      for (int64_t x = box.lo.x + blockIdx.x * blockDim.x + threadIdx.x; x < box.hi.x; x += gridDim.x * blockDim.x) {
        const Dim3 p(x, y, z);
        dst[p] = field(p);
      }
}

/* next = (d/dx + d/dy + d/dz) curr by centred differences */
__global__ void gradient_sum(Accessor<float> dst, Accessor<float> src, Rect3 box) {
  for (int64_t z = box.lo.z + blockIdx.z; z < box.hi.z; z += gridDim.z)
    for (int64_t y = box.lo.y + blockIdx.y * blockDim.y + threadIdx.y; y < box.hi.y; y += gridDim.y * blockDim.y)
      for (int64_t x = box.lo.x + blockIdx.x * blockDim.x + threadIdx.x; x < box.hi.x; x += gridDim.x * blockDim.x) {
        const Dim3 p(x, y, z);
        const float dx = (src[p + Dim3(1, 0, 0)] - src[p - Dim3(1, 0, 0)]) * 0.5f;
        const float dy = (src[p + Dim3(0, 1, 0)] - src[p - Dim3(0, 1, 0)]) * 0.5f;
        const float dz = (src[p + Dim3(0, 0, 1)] - src[p - Dim3(0, 0, 1)]) * 0.5f;
        dst[p] = dx + dy + dz;
      }
}

float expect_axis(int64_t v, int64_t n) {
  /* centred difference of t + ripple(t) with periodic wrap of the index (values jump by n at the seam) */
  const int64_t hi = (v + 1) % n, lo = (v - 1 + n) % n;
  return ((float(hi) + ripple_at(hi)) - (float(lo) + ripple_at(lo))) * 0.5f;
}

size_t run_case(const Dim3 &size, const std::vector<int> &gpus, MethodFlags methods) {
  DistributedDomain dd(size_t(size.x), size_t(size.y), size_t(size.z));
  dd.set_radius(1);
  auto h = dd.add_data<float>("f");
  dd.set_methods(methods);
  dd.set_gpus(gpus);
  dd.set_plan_file(false);
  dd.realize();

  const dim3 block(32, 4, 1), grid(2, 4, 8);
  for (LocalDomain &d : dd.domains()) {
    CUDA_RUNTIME(cudaSetDevice(d.gpu()));
    ripple_init<<<grid, block>>>(d.get_curr_accessor(h), d.get_compute_region());
    CUDA_RUNTIME(cudaDeviceSynchronize());
  }

  /* interior first (needs no halo), then the exchange, then the exterior slabs — the reference's overlap order */
  const std::vector<Rect3> interiors = dd.get_interior();
  const std::vector<std::vector<Rect3>> exteriors = dd.get_exterior();
  for (size_t di = 0; di < dd.domains().size(); ++di) {
    LocalDomain &d = dd.domains()[di];
    CUDA_RUNTIME(cudaSetDevice(d.gpu()));
    gradient_sum<<<grid, block>>>(d.get_next_accessor(h), d.get_curr_accessor(h), interiors[di]);
  }
  dd.exchange();
  for (size_t di = 0; di < dd.domains().size(); ++di) {
    LocalDomain &d = dd.domains()[di];
    CUDA_RUNTIME(cudaSetDevice(d.gpu()));
    for (const Rect3 &slab : exteriors[di])
      if (slab.extent().flatten() > 0) gradient_sum<<<grid, block>>>(d.get_next_accessor(h), d.get_curr_accessor(h), slab);
    CUDA_RUNTIME(cudaDeviceSynchronize());
  }
  dd.swap();

  size_t wrong = 0;
  for (LocalDomain &d : dd.domains()) {
    const std::vector<unsigned char> bytes = d.interior_to_host(h.id());
    std::vector<float> v(bytes.size() / sizeof(float));
    std::memcpy(v.data(), bytes.data(), bytes.size());
    const Dim3 sz = d.size(), org = d.origin();
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x, ++i) {
          const Dim3 p = org + Dim3(x, y, z);
          const float want = expect_axis(p.x, size.x) + expect_axis(p.y, size.y) + expect_axis(p.z, size.z);
          if (v[i] != want) ++wrong;
        }
  }
  return wrong;
}

} // namespace

TEST_CASE("derivative: centred differences across subdomain and periodic boundaries", "[cuda][derivative]") {
  /* sizes are multiples of 4 so the ripple is periodic with the domain; all sums are exact in fp32 */
  REQUIRE(run_case(Dim3(12, 8, 16), {0}, MethodFlags::All) == 0);
  REQUIRE(run_case(Dim3(12, 8, 16), {0}, MethodFlags::CudaMpi) == 0); // the reference's configuration: staged only
  REQUIRE(run_case(Dim3(24, 12, 8), {0, 0}, MethodFlags::All) == 0);
  REQUIRE(run_case(Dim3(24, 12, 8), {0, 0, 0}, MethodFlags::CudaKernel | MethodFlags::CudaMpi) == 0);
  REQUIRE(run_case(Dim3(20, 16, 12), {0, 0, 0, 0}, MethodFlags::CudaMemcpyPeer | MethodFlags::CudaMpi) == 0);
}

TEST_CASE("derivative: every visible GPU", "[cuda][derivative][multigpu]") {
  int n = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&n));
  if (n < 2) return;
  std::vector<int> gpus;
  for (int i = 0; i < n; ++i) gpus.push_back(i);
  REQUIRE(run_case(Dim3(64, 32, 24), gpus, MethodFlags::All) == 0);
  REQUIRE(run_case(Dim3(64, 32, 24), gpus, MethodFlags::CudaMemcpyPeer | MethodFlags::CudaMpi) == 0);
}
