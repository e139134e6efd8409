/* The smallest complete program on the library: a 7-point heat stencil written by the USER (not the Jacobi3D engine),
   on every visible GPU, with the interior / exchange / exterior overlap the reference's README teaches and — when the
   plan needs no host (one NVSwitch box) — without ever blocking the host on the exchange.

     ./build/bin/example-heat [nx ny nz iters]

   Prints the mean temperature, which a periodic average-of-neighbours stencil conserves. */
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "stencil/stencil.hpp"

__global__ void init(Accessor<float> dst, Rect3 box) {
  for (int64_t z = box.lo.z + blockIdx.z; z < box.hi.z; z += gridDim.z)
    for (int64_t y = box.lo.y + blockIdx.y * blockDim.y + threadIdx.y; y < box.hi.y; y += gridDim.y * blockDim.y)
      for (int64_t x = box.lo.x + blockIdx.x * blockDim.x + threadIdx.x; x < box.hi.x; x += gridDim.x * blockDim.x)
        dst[Dim3(x, y, z)] = ((x / 8 + y / 8 + z / 8) & 1) ? 1.f : 0.f; // a checkerboard of hot and cold blocks
}

__global__ void relax(Accessor<float> dst, Accessor<float> src, Rect3 box) {
  for (int64_t z = box.lo.z + blockIdx.z; z < box.hi.z; z += gridDim.z)
    for (int64_t y = box.lo.y + blockIdx.y * blockDim.y + threadIdx.y; y < box.hi.y; y += gridDim.y * blockDim.y)
      for (int64_t x = box.lo.x + blockIdx.x * blockDim.x + threadIdx.x; x < box.hi.x; x += gridDim.x * blockDim.x) {
        const Dim3 p(x, y, z);
        dst[p] = (src[p + Dim3(1, 0, 0)] + src[p - Dim3(1, 0, 0)] + src[p + Dim3(0, 1, 0)] + src[p - Dim3(0, 1, 0)] + src[p + Dim3(0, 0, 1)] +
                  src[p - Dim3(0, 0, 1)]) /
                 6.f;
      }
}

int main(int argc, char **argv) {
  const size_t nx = argc > 1 ? size_t(std::atoi(argv[1])) : 128, ny = argc > 2 ? size_t(std::atoi(argv[2])) : 128,
               nz = argc > 3 ? size_t(std::atoi(argv[3])) : 128;
  const int iters = argc > 4 ? std::atoi(argv[4]) : 20;

  DistributedDomain dd(nx, ny, nz);
  Radius radius = Radius::constant(0);
  radius.set_face(1); // a 7-point stencil reads faces only
  dd.set_radius(radius);
  const DataHandle<float> temp = dd.add_data<float>("temperature");
  dd.set_methods(MethodFlags::All);
  dd.set_plan_file(false);
  dd.realize();

  const dim3 block(32, 8, 1), grid(4, 8, 16);
  std::vector<LocalDomain> &doms = dd.domains();
  std::vector<RcStream> streams;
  std::vector<cudaEvent_t> done(doms.size());
  for (size_t i = 0; i < doms.size(); ++i) {
    CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
    streams.emplace_back(doms[i].gpu());
    CUDA_RUNTIME(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
    init<<<grid, block, 0, streams[i]>>>(doms[i].get_curr_accessor(temp), doms[i].get_compute_region());
    CUDA_RUNTIME(cudaEventRecord(done[i], streams[i]));
  }

  const std::vector<Rect3> interior = dd.get_interior();
  const std::vector<std::vector<Rect3>> exterior = dd.get_exterior();
  const bool hostFree = dd.exchange_device_only();
  for (int it = 0; it < iters; ++it) {
    /* 1. the exchange starts as soon as the kernels that wrote `curr` (and read the old halos) are done */
    for (size_t i = 0; i < doms.size(); ++i) dd.exchange_after(i, done[i]);
    dd.exchange_async();
    /* 2. interior cells need no halo: they overlap the exchange */
    for (size_t i = 0; i < doms.size(); ++i) {
      CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
      relax<<<grid, block, 0, streams[i]>>>(doms[i].get_next_accessor(temp), doms[i].get_curr_accessor(temp), interior[i]);
    }
    /* 3. exterior slabs wait for the halos: by an event when no transport needs the host, else through exchange_wait() */
    if (!hostFree) dd.exchange_wait();
    for (size_t i = 0; i < doms.size(); ++i) {
      CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
      if (hostFree) CUDA_RUNTIME(cudaStreamWaitEvent(streams[i], dd.exchange_done_event(i), 0));
      for (const Rect3 &slab : exterior[i])
        if (slab.extent().flatten() > 0) relax<<<grid, block, 0, streams[i]>>>(doms[i].get_next_accessor(temp), doms[i].get_curr_accessor(temp), slab);
      CUDA_RUNTIME(cudaEventRecord(done[i], streams[i]));
    }
    dd.swap();
  }
  for (size_t i = 0; i < doms.size(); ++i) {
    CUDA_RUNTIME(cudaSetDevice(doms[i].gpu()));
    CUDA_RUNTIME(cudaStreamSynchronize(streams[i]));
  }
  dd.exchange_wait();

  double sum = 0;
  for (LocalDomain &d : doms) {
    const std::vector<unsigned char> bytes = d.interior_to_host(temp.id());
    const float *v = reinterpret_cast<const float *>(bytes.data());
    for (size_t k = 0; k < bytes.size() / sizeof(float); ++k) sum += v[k];
  }
  sum = double(dd.comm().allreduce_sum(uint64_t(sum + 0.5))); // whole cells of heat: exact enough for a demo
  if (dd.rank() == 0) std::printf("mean temperature after %d steps: %.6f\n", iters, sum / (double(nx) * double(ny) * double(nz)));
  for (cudaEvent_t e : done) cudaEventDestroy(e);
  return 0;
}
