/* bench-pack: pack and unpack time of one face message on a 512^3 radius-3 fp32 LocalDomain, CUDA-event timed.
   CLI and output follow the reference's bin/bench_pack.cu:94-107:  --iters (30);
   one line per direction (+x, +y, +z):  <ext> <dir> <bytes> <pack s> <unpack s>
   Additions: --x --y --z --r --q to change the domain, radius and number of fp32 quantities. */
#include <iostream>

#include "stencil/argparse.hpp"
#include "stencil/packer.cuh"
#include "stencil/rcstream.hpp"

int main(int argc, char **argv) {
  int nIters = 30;
  size_t x = 512, y = 512, z = 512, r = 3, q = 1;
  argparse::Parser p("pack/unpack kernel benchmark");
  p.add_option(nIters, "--iters");
  p.add_option(x, "--x");
  p.add_option(y, "--y");
  p.add_option(z, "--z");
  p.add_option(r, "--r");
  p.add_option(q, "--q");
  if (!p.parse(argc, argv) || p.need_help()) {
    std::cout << p.help();
    return p.need_help() ? 0 : EXIT_FAILURE;
  }
  CUDA_RUNTIME(cudaSetDevice(0));
  RcStream stream(0);
  LocalDomain ld(Dim3(int64_t(x), int64_t(y), int64_t(z)), Dim3(0, 0, 0), 0);
  ld.set_radius(r);
  for (size_t i = 0; i < q; ++i) ld.add_data<float>();
  ld.realize();
  cudaEvent_t t0, t1;
  CUDA_RUNTIME(cudaEventCreate(&t0));
  CUDA_RUNTIME(cudaEventCreate(&t1));
  for (const Dim3 dir : {Dim3(1, 0, 0), Dim3(0, 1, 0), Dim3(0, 0, 1)}) {
    std::vector<Message> msgs = {Message(dir, 0, 0)};
    DevicePacker packer(stream);
    DeviceUnpacker unpacker(stream);
    packer.prepare(&ld, msgs);
    unpacker.prepare(&ld, msgs);
    float packMs = 0, unpackMs = 0;
    packer.pack(); // warm up
    unpacker.unpack();
    CUDA_RUNTIME(cudaEventRecord(t0, stream));
    for (int i = 0; i < nIters; ++i) packer.pack();
    CUDA_RUNTIME(cudaEventRecord(t1, stream));
    CUDA_RUNTIME(cudaEventSynchronize(t1));
    CUDA_RUNTIME(cudaEventElapsedTime(&packMs, t0, t1));
    CUDA_RUNTIME(cudaEventRecord(t0, stream));
    for (int i = 0; i < nIters; ++i) unpacker.unpack();
    CUDA_RUNTIME(cudaEventRecord(t1, stream));
    CUDA_RUNTIME(cudaEventSynchronize(t1));
    CUDA_RUNTIME(cudaEventElapsedTime(&unpackMs, t0, t1));
    std::cout << ld.halo_extent(dir * -1) << " " << dir << " " << packer.size() << " " << double(packMs) / 1e3 / nIters << " "
              << double(unpackMs) / 1e3 / nIters << "\n";
  }
  cudaEventDestroy(t0);
  cudaEventDestroy(t1);
  return 0;
}
