/* measure-buf-exchange: find, for every GPU pair, the message size that takes a fixed time (4 ms +-1 %) when ALL pairs
   copy at once — the effective bandwidth matrix under full contention.
   Reference: bin/measure_buf_exchange.cu (hard-coded to 4 GPUs). Same method, any GPU count (--gpus N): start at
   8 MiB per pair, run all cudaMemcpyPeerAsync copies together behind a clock_block latch so they start at the same
   moment, measure each with its own events, move every size by gamma = 0.2 of its relative error, repeat. Prints the
   size matrix `x` (MiB), the time matrix `y` (s) and the correction `dx` each round, then the final MiB matrix and
   the implied GB/s matrix (which can be fed to gpu_topo::set_bandwidth_matrix / the NodeAware placement). */
#include <cmath>
#include <cstdio>
#include <iostream>

#include "stencil/argparse.hpp"
#include "stencil/cuda_runtime.hpp"
#include "stencil/gpu_topology.hpp"
#include "stencil/mat2d.hpp"

/* spin for `cycles` clock ticks: queued copies on other streams start together when this kernel ends */
__global__ void clock_block(long long cycles) {
  const long long start = clock64();
  while (clock64() - start < cycles) {
  }
}

int main(int argc, char **argv) {
  int n = 0, rounds = 50, measures = 10;
  double period = 0.004, gamma = 0.2;
  argparse::Parser p("equal-time message sizes for all GPU pairs");
  p.add_option(n, "--gpus");
  p.add_option(rounds, "--rounds");
  p.add_option(measures, "--measures");
  p.add_option(period, "--period");
  if (!p.parse(argc, argv) || p.need_help()) {
    std::cout << p.help();
    return p.need_help() ? 0 : EXIT_FAILURE;
  }
  int devCount = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&devCount));
  if (n <= 0 || n > devCount) n = devCount;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) gpu_topo::enable_peer(i, j);
  const int64_t M = 1024 * 1024;
  Mat2D<cudaStream_t> streams(n, n, nullptr);
  Mat2D<cudaEvent_t> startEv(n, n, nullptr), stopEv(n, n, nullptr);
  Mat2D<void *> src(n, n, nullptr), dst(n, n, nullptr);
  Mat2D<int64_t> cap(n, n, 0), x(n, n, 8 * M), dx(n, n, 0);
  Mat2D<double> y(n, n, 0.0);
  cudaEvent_t latch;
  CUDA_RUNTIME(cudaSetDevice(0));
  CUDA_RUNTIME(cudaEventCreate(&latch));
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      CUDA_RUNTIME(cudaSetDevice(i));
      CUDA_RUNTIME(cudaStreamCreate(&streams.at(i, j)));
      CUDA_RUNTIME(cudaEventCreate(&startEv.at(i, j)));
      CUDA_RUNTIME(cudaEventCreate(&stopEv.at(i, j)));
    }
  auto print = [&](const char *name, auto get) {
    std::cout << name << "\n";
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) std::cout << get(i, j) << " ";
      std::cout << "\n";
    }
  };
  for (int round = 0; round < rounds; ++round) {
    print("x", [&](int i, int j) { return x.at(i, j) / M; });
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j)
        if (cap.at(i, j) < x.at(i, j)) {
          cap.at(i, j) = x.at(i, j) * 2;
          CUDA_RUNTIME(cudaSetDevice(i));
          CUDA_RUNTIME(cudaFree(src.at(i, j)));
          CUDA_RUNTIME(cudaMalloc(&src.at(i, j), size_t(cap.at(i, j))));
          CUDA_RUNTIME(cudaSetDevice(j));
          CUDA_RUNTIME(cudaFree(dst.at(i, j)));
          CUDA_RUNTIME(cudaMalloc(&dst.at(i, j), size_t(cap.at(i, j))));
        }
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) y.at(i, j) = 0;
    for (int m = 0; m < measures; ++m) {
      CUDA_RUNTIME(cudaSetDevice(0));
      clock_block<<<1, 1>>>(2000000); // ~1 ms at 2 GHz: long enough to enqueue every copy behind it
      CUDA_RUNTIME(cudaEventRecord(latch, 0));
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
          CUDA_RUNTIME(cudaSetDevice(i));
          CUDA_RUNTIME(cudaStreamWaitEvent(streams.at(i, j), latch, 0));
          CUDA_RUNTIME(cudaEventRecord(startEv.at(i, j), streams.at(i, j)));
          CUDA_RUNTIME(cudaMemcpyPeerAsync(dst.at(i, j), j, src.at(i, j), i, size_t(x.at(i, j)), streams.at(i, j)));
          CUDA_RUNTIME(cudaEventRecord(stopEv.at(i, j), streams.at(i, j)));
        }
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
          CUDA_RUNTIME(cudaSetDevice(i));
          CUDA_RUNTIME(cudaEventSynchronize(stopEv.at(i, j)));
          float ms = 0;
          CUDA_RUNTIME(cudaEventElapsedTime(&ms, startEv.at(i, j), stopEv.at(i, j)));
          y.at(i, j) += double(ms) / 1e3 / measures;
        }
    }
    print("y", [&](int i, int j) { return y.at(i, j); });
    bool converged = true;
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
        const double err = (period - y.at(i, j)) / period;
        if (std::fabs(err) > 0.01) converged = false;
        dx.at(i, j) = int64_t(gamma * err * double(x.at(i, j)));
        x.at(i, j) = std::max<int64_t>(M / 16, x.at(i, j) + dx.at(i, j));
      }
    print("dx", [&](int i, int j) { return dx.at(i, j); });
    if (converged) break;
  }
  print("final MiB", [&](int i, int j) { return double(x.at(i, j)) / double(M); });
  print("GB/s", [&](int i, int j) { return double(x.at(i, j)) / y.at(i, j) / 1e9; });
  return 0;
}
