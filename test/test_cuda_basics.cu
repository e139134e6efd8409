/* Accessor on device, allocators, Array<gpu>, gpu_topo, RcStream.
   Same ground as the reference's test_cuda_{accessor,allocator,array,gpu_topo,rcstream,align}.cu. */
#include "harness.hpp"
#include <vector>
#include "stencil/accessor.hpp"
#include "stencil/align.cuh"
#include "stencil/array.hpp"
#include "stencil/device_allocator.hpp"
#include "stencil/gpu_topology.hpp"
#include "stencil/managed_allocator.hpp"
#include "stencil/rcstream.hpp"

__global__ void fill_via_accessor(Accessor<int> acc, Dim3 lo, Dim3 hi) {
  for (int64_t z = lo.z + blockIdx.z; z < hi.z; z += gridDim.z)
    for (int64_t y = lo.y + threadIdx.y; y < hi.y; y += blockDim.y)
      for (int64_t x = lo.x + threadIdx.x; x < hi.x; x += blockDim.x) acc[Dim3(x, y, z)] = int(x + 100 * y + 10000 * z);
}

__global__ void align_on_device(size_t *out) { out[0] = next_align_of(13, 8); }

TEST_CASE("accessor indexes with global coordinates and a padded pitch", "[cuda][accessor]") {
  const Dim3 origin(5, 6, 7), ext(3, 4, 2), pitch(8, 4, 2); // row pitch 8 > extent 3
  std::vector<int, ManagedAllocator<int>> buf(size_t(pitch.flatten()), -1, ManagedAllocator<int>(0));
  Accessor<int> acc(buf.data(), origin, pitch);
  fill_via_accessor<<<dim3(1, 1, 2), dim3(4, 4, 1)>>>(acc, origin, origin + ext);
  CUDA_RUNTIME(cudaDeviceSynchronize());
  for (int64_t z = 0; z < ext.z; ++z)
    for (int64_t y = 0; y < ext.y; ++y)
      for (int64_t x = 0; x < pitch.x; ++x) {
        const int v = buf[size_t((z * pitch.y + y) * pitch.x + x)];
        if (x < ext.x)
          REQUIRE(v == int((origin.x + x) + 100 * (origin.y + y) + 10000 * (origin.z + z)));
        else
          REQUIRE(v == -1); // padding untouched
      }
  REQUIRE(acc[Dim3(6, 7, 8)] == 6 + 700 + 80000); // host side of the same accessor (managed memory)
  REQUIRE(acc.origin() == origin);
  REQUIRE(acc.pitch() == pitch);
}

TEST_CASE("next_align_of on device", "[cuda][align]") {
  size_t *d = nullptr, h = 0;
  CUDA_RUNTIME(cudaMalloc(&d, sizeof(size_t)));
  align_on_device<<<1, 1>>>(d);
  CUDA_RUNTIME(cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost));
  CUDA_RUNTIME(cudaFree(d));
  REQUIRE(h == 16);
}

TEST_CASE("device and managed allocators", "[cuda][allocator]") {
  DeviceAllocator<double> da(0);
  double *p = da.allocate(100);
  REQUIRE(p != nullptr);
  cudaPointerAttributes attr;
  CUDA_RUNTIME(cudaPointerGetAttributes(&attr, p));
  REQUIRE(attr.type == cudaMemoryTypeDevice);
  REQUIRE(attr.device == 0);
  da.deallocate(p, 100);
  REQUIRE(da.max_size() == size_t(-1) / sizeof(double));
  REQUIRE(ManagedAllocator<float>(0).max_size() == size_t(-1) / sizeof(float));
  std::vector<float, ManagedAllocator<float>> v(10, 1.5f, ManagedAllocator<float>(0));
  CUDA_RUNTIME(cudaPointerGetAttributes(&attr, v.data()));
  REQUIRE(attr.type == cudaMemoryTypeManaged);
  REQUIRE(v[9] == 1.5f);
}

TEST_CASE("array on the gpu", "[cuda][array]") {
  Array<int, device_type::gpu> a(8);
  REQUIRE(a.size() == 8);
  int h[8] = {0, 1, 2, 3, 4, 5, 6, 7}, back[16] = {0};
  a.from_host(h, 8);
  a.resize(16); // keeps the first 8
  REQUIRE(a.size() == 16);
  a.to_host(back, 8);
  for (int i = 0; i < 8; ++i) REQUIRE(back[i] == i);
  cudaPointerAttributes attr;
  CUDA_RUNTIME(cudaPointerGetAttributes(&attr, a.data()));
  REQUIRE(attr.type == cudaMemoryTypeDevice);
}

TEST_CASE("gpu_topo self and pairs", "[cuda][topo]") {
  gpu_topo::enable_peer(0, 0);
  REQUIRE(gpu_topo::peer(0, 0));
  REQUIRE(!gpu_topo::peer(0, 63)); // never registered
  REQUIRE(gpu_topo::bandwidth(0, 0) == 10.0);
  int n = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&n));
  for (int j = 1; j < n; ++j) {
    REQUIRE(gpu_topo::bandwidth(0, j) > 0);
    REQUIRE(gpu_topo::bandwidth(0, j) < gpu_topo::bandwidth(0, 0));
    gpu_topo::enable_peer(0, j);
  }
  Mat2D<double> fake = {{9, 2}, {2, 9}};
  gpu_topo::set_bandwidth_matrix(fake);
  REQUIRE(gpu_topo::bandwidth(0, 1) == 2);
  gpu_topo::set_bandwidth_matrix(Mat2D<double>());
  REQUIRE(gpu_topo::bandwidth(0, 0) == 10.0);
}

TEST_CASE("rcstream sharing and lifetime", "[cuda][rcstream]") {
  RcStream a(0);
  REQUIRE(cudaStream_t(a) != nullptr);
  REQUIRE(a.device() == 0);
  REQUIRE(a.use_count() == 1);
  {
    RcStream b(a); // copy shares
    REQUIRE(b == a);
    REQUIRE(a.use_count() == 2);
    RcStream c(std::move(b)); // move transfers
    REQUIRE(c == a);
    REQUIRE(a.use_count() == 2);
  }
  REQUIRE(a.use_count() == 1);
  RcStream hi(0, RcStream::Priority::HIGH);
  int prio = 0, least = 0, greatest = 0;
  CUDA_RUNTIME(cudaStreamGetPriority(hi, &prio));
  CUDA_RUNTIME(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  REQUIRE(prio == greatest);
  REQUIRE(hi != a);
  RcStream d;
  d = hi; // copy assignment releases the old stream
  REQUIRE(d == hi);
  CUDA_RUNTIME(cudaStreamSynchronize(d));
}
