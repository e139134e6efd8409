/* bench-alltoallv: all-pairs GPU-to-GPU traffic matrices inside one process, mean wall time per exchange.
   Reference: bin/bench_alltoallv.cu (hard-coded to 4 GPUs, cudaMemcpyPeerAsync only). Here the five cases are
   generated for however many GPUs are visible (or --gpus N), and each case is run twice: with cudaMemcpyPeerAsync on
   per-pair streams (the reference's method) and with device-initiated stores (the library's box-copy kernel writing
   through peer pointers), which is what the NvlinkDirect exchange uses. Output per case: the reference's two lines
   (label, seconds for cudaMemcpyPeer) plus a third line "  direct-store <seconds>".
   Options: --gpus N, --iters (30), --cap-mib M (largest single message, default 1024). */
#include <chrono>
#include <iostream>

#include "stencil/argparse.hpp"
#include "stencil/copy_plan.hpp"
#include "stencil/cuda_runtime.hpp"
#include "stencil/gpu_topology.hpp"
#include "stencil/mat2d.hpp"
#include "stencil/partition.hpp"

typedef std::chrono::steady_clock Clock;

struct Pairwise {
  int n;
  Mat2D<int64_t> bytes;
  std::vector<void *> src, dst; // [i*n+j]: src on i, dst on j
  std::vector<cudaStream_t> streams;
  std::vector<stencil::CopyPlan> plans; // per source GPU: all its outgoing messages in one launch
  std::vector<cudaStream_t> planStreams;

  explicit Pairwise(const Mat2D<int64_t> &m) : n(int(m.shape().x)), bytes(m), src(size_t(n * n)), dst(size_t(n * n)), streams(size_t(n * n)), plans(size_t(n)), planStreams(size_t(n)) {
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) {
        const size_t k = size_t(i * n + j), b = size_t(bytes.at(i, j));
        CUDA_RUNTIME(cudaSetDevice(i));
        CUDA_RUNTIME(cudaStreamCreate(&streams[k]));
        if (!b) continue;
        CUDA_RUNTIME(cudaMalloc(&src[k], b));
        CUDA_RUNTIME(cudaSetDevice(j));
        CUDA_RUNTIME(cudaMalloc(&dst[k], b));
      }
    }
    for (int i = 0; i < n; ++i) {
      CUDA_RUNTIME(cudaSetDevice(i));
      CUDA_RUNTIME(cudaStreamCreate(&planStreams[size_t(i)]));
      for (int j = 0; j < n; ++j) {
        const size_t k = size_t(i * n + j);
        const int64_t b = bytes.at(i, j);
        if (!b) continue;
        /* a message is a box of 64 KiB rows (+ one shorter row for the rest): the plan spreads ROWS over CTAs, so one
           long row would be copied by a single warp */
        const int64_t units = b / 16, rowUnits = 4096, rows = units / rowUnits, rest = units % rowUnits;
        if (rows) {
          const Dim3 ext(rowUnits, rows, 1);
          plans[size_t(i)].add(stencil::BoxView::dense(src[k], ext, 16), Dim3(0, 0, 0), stencil::BoxView::dense(dst[k], ext, 16), Dim3(0, 0, 0), ext);
        }
        if (rest) {
          const Dim3 ext(rest, 1, 1);
          char *so = static_cast<char *>(src[k]) + rows * rowUnits * 16, *dn = static_cast<char *>(dst[k]) + rows * rowUnits * 16;
          plans[size_t(i)].add(stencil::BoxView::dense(so, ext, 16), Dim3(0, 0, 0), stencil::BoxView::dense(dn, ext, 16), Dim3(0, 0, 0), ext);
        }
      }
      plans[size_t(i)].finalize(i, 256 * 1024);
    }
  }
  ~Pairwise() {
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
        const size_t k = size_t(i * n + j);
        cudaSetDevice(i);
        cudaFree(src[k]);
        cudaStreamDestroy(streams[k]);
        cudaSetDevice(j);
        cudaFree(dst[k]);
      }
    for (int i = 0; i < n; ++i) {
      cudaSetDevice(i);
      cudaStreamDestroy(planStreams[size_t(i)]);
    }
  }
  void sync_all() {
    for (int i = 0; i < n; ++i) {
      CUDA_RUNTIME(cudaSetDevice(i));
      CUDA_RUNTIME(cudaDeviceSynchronize());
    }
  }
  double run_memcpy_peer(int iters) {
    sync_all();
    const auto t0 = Clock::now();
    for (int it = 0; it < iters; ++it) {
      for (int i = 0; i < n; ++i) {
        CUDA_RUNTIME(cudaSetDevice(i));
        for (int j = 0; j < n; ++j) {
          const size_t k = size_t(i * n + j);
          if (bytes.at(i, j)) CUDA_RUNTIME(cudaMemcpyPeerAsync(dst[k], j, src[k], i, size_t(bytes.at(i, j)), streams[k]));
        }
      }
      sync_all();
    }
    return std::chrono::duration<double>(Clock::now() - t0).count() / iters;
  }
  double run_direct(int iters) {
    sync_all();
    const auto t0 = Clock::now();
    for (int it = 0; it < iters; ++it) {
      for (int i = 0; i < n; ++i) {
        CUDA_RUNTIME(cudaSetDevice(i));
        plans[size_t(i)].launch(0, planStreams[size_t(i)]);
      }
      sync_all();
    }
    return std::chrono::duration<double>(Clock::now() - t0).count() / iters;
  }
};

int main(int argc, char **argv) {
  int n = 0, iters = 30;
  int64_t capMiB = 1024;
  argparse::Parser p("all-pairs GPU traffic benchmark");
  p.add_option(n, "--gpus");
  p.add_option(iters, "--iters");
  p.add_option(capMiB, "--cap-mib");
  if (!p.parse(argc, argv) || p.need_help()) {
    std::cout << p.help();
    return p.need_help() ? 0 : EXIT_FAILURE;
  }
  int devCount = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&devCount));
  if (n <= 0 || n > devCount) n = devCount;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) gpu_topo::enable_peer(i, j);
  const int64_t K = 1024, M = K * K, G = K * M, cap = capMiB * M;
  auto capped = [&](int64_t v) { return std::min(v, cap) / 16 * 16; };
  auto filled = [&](int64_t v) { return Mat2D<int64_t>(n, n, capped(v)); };

  /* "stencil": bytes a 512x256x512-per-GPU, radius-3, 4 x fp32 decomposition moves between each pair (incl. self) */
  Mat2D<int64_t> stencil(n, n, 0);
  {
    const Radius r = Radius::constant(3);
    NodePartition part(Dim3(512, 256, 512) * Dim3(1, 1, 1), r, 1, n);
    NodePartition whole(Dim3(512, 256, 512) * part.dim(), r, 1, n);
    const Mat2D<double> cells = NodeAware::comm_matrix(whole, Dim3(0, 0, 0), r);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) stencil.at(i, j) = capped(int64_t(cells.at(i, j)) * 16);
  }
  Mat2D<int64_t> local = filled(0), localRemote = filled(100 * M);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j)
      if (i / std::max(1, n / 2) == j / std::max(1, n / 2)) {
        local.at(i, j) = capped(G);
        localRemote.at(i, j) = capped(G);
      }
  const struct {
    const char *label;
    Mat2D<int64_t> m;
  } cases[] = {{"stencil", stencil}, {"All-to-all 8MiB", filled(8 * M)}, {"All-to-all 1GiB", filled(G)}, {"Local 1GiB", local},
               {"Local 1GiB Remote 100M", localRemote}};
  for (const auto &c : cases) {
    Pairwise pw(c.m);
    pw.run_memcpy_peer(2);
    pw.run_direct(2);
    std::cout << c.label << "\n" << pw.run_memcpy_peer(iters) << "\n";
    std::cout << "  direct-store " << pw.run_direct(iters) << "\n";
  }
  return 0;
}
