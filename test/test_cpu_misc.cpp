/* Statistics, argparse, align, halo geometry, Comm (self + TCP loopback), MpiTopology. */
#include "harness.hpp"
#include <cmath>
#include <thread>
#include <unistd.h>
#include "stencil/boundary.hpp"
#include "stencil/align.cuh"
#include "stencil/argparse.hpp"
#include "stencil/comm.hpp"
#include "stencil/halo_geometry.hpp"
#include "stencil/mpi_topology.hpp"
#include "stencil/statistics.hpp"

TEST_CASE("statistics", "[stats]") {
  Statistics s;
  REQUIRE(std::isnan(s.min()));
  REQUIRE(std::isnan(s.trimean()));
  for (double v : {4.0, 1.0, 3.0, 2.0, 5.0, 8.0, 7.0, 6.0}) s.insert(v);
  REQUIRE(s.count() == 8);
  REQUIRE(s.min() == 1);
  REQUIRE(s.max() == 8);
  REQUIRE(s.avg() == 4.5);
  REQUIRE(s.med() == 4.5);
  REQUIRE(s.trimean() == (3 + 2 * 5 + 7) / 4.0); // sorted x[2], x[4], x[6]
  REQUIRE(std::abs(s.stddev() - std::sqrt(6.0)) < 1e-12);
  s.insert(9);
  REQUIRE(s.med() == 5);
  s.clear();
  REQUIRE(s.count() == 0);
}

TEST_CASE("argparse", "[argparse]") {
  bool flag = false, other = false;
  int iters = 0;
  std::string prefix;
  size_t x = 0, y = 0;
  double f = 0;
  argparse::Parser p("test");
  p.add_flag(flag, "--flag", "-f")->help("a flag");
  p.add_flag(other, "--other");
  p.add_option(iters, "--iters", "-n");
  p.add_option(prefix, "--prefix");
  p.add_option(f, "--scale");
  p.add_positional(x)->required();
  p.add_positional(y);
  const char *args[] = {"prog", "12", "-n", "30", "--flag", "--prefix", "out_", "--scale", "-1.5", "34", "--unknown"};
  int argc = 11;
  char *argv[11];
  for (int i = 0; i < argc; ++i) argv[i] = const_cast<char *>(args[i]);
  REQUIRE(p.parse(argc, argv));
  REQUIRE(flag);
  REQUIRE(!other);
  REQUIRE(iters == 30);
  REQUIRE(prefix == "out_");
  REQUIRE(f == -1.5);
  REQUIRE(x == 12);
  REQUIRE(y == 34);
  REQUIRE(argc == 2); // prog + the unrecognised one are left
  REQUIRE(std::string(argv[1]) == "--unknown");
  REQUIRE(!p.need_help());
  REQUIRE(p.help().find("--iters") != std::string::npos);

  argparse::Parser q;
  size_t need = 0;
  q.add_positional(need)->required();
  int c1 = 1;
  char *a1[1] = {const_cast<char *>("prog")};
  REQUIRE(!q.parse(c1, a1)); // missing required positional

  argparse::Parser r;
  r.no_unrecognized();
  int c2 = 2;
  char *a2[2] = {const_cast<char *>("prog"), const_cast<char *>("--nope")};
  REQUIRE(!r.parse(c2, a2));
}

TEST_CASE("align", "[align]") {
  REQUIRE(next_align_of(0, 8) == 0);
  REQUIRE(next_align_of(1, 8) == 8);
  REQUIRE(next_align_of(8, 8) == 8);
  REQUIRE(next_align_of(9, 4) == 12);
  REQUIRE(next_align_of(100, 1) == 100);
  REQUIRE(round_up_multiple(10, 6) == 12);
}

TEST_CASE("halo geometry", "[geometry]") {
  Radius r = Radius::constant(4);
  const Dim3 sz(30, 40, 50);
  REQUIRE(stencil::halo_extent(Dim3(1, 0, 0), sz, r) == Dim3(4, 40, 50));
  REQUIRE(stencil::halo_extent(Dim3(0, -1, 1), sz, r) == Dim3(30, 4, 4));
  REQUIRE(stencil::halo_extent(Dim3(0, 0, 0), sz, r) == sz);
  REQUIRE(stencil::halo_pos(Dim3(1, 0, 0), sz, r, true) == Dim3(34, 4, 4));
  REQUIRE(stencil::halo_pos(Dim3(1, 0, 0), sz, r, false) == Dim3(30, 4, 4));
  REQUIRE(stencil::halo_pos(Dim3(-1, 0, 0), sz, r, true) == Dim3(0, 4, 4));
  REQUIRE(stencil::halo_pos(Dim3(-1, 0, 0), sz, r, false) == Dim3(4, 4, 4));
  REQUIRE(stencil::halo_pos(Dim3(0, 0, 0), sz, r, true) == Dim3(4, 4, 4));
  Radius u = Radius::constant(0);
  u.dir(1, 0, 0) = 2;
  u.dir(-1, 0, 0) = 1;
  REQUIRE(stencil::halo_extent(Dim3(1, 0, 0), sz, u) == Dim3(2, 40, 50));
  REQUIRE(stencil::halo_extent(Dim3(-1, 0, 0), sz, u) == Dim3(1, 40, 50));
  REQUIRE(stencil::halo_pos(Dim3(1, 0, 0), sz, u, true) == Dim3(31, 0, 0));
  REQUIRE(stencil::halo_pos(Dim3(1, 0, 0), sz, u, false) == Dim3(30, 0, 0)); // last r(-x)=1 compute cells
}

TEST_CASE("self comm", "[comm]") {
  std::unique_ptr<stencil::Comm> c = stencil::make_self_comm();
  REQUIRE(c->rank() == 0);
  REQUIRE(c->size() == 1);
  int v = 42, w = 0;
  REQUIRE(!c->try_recv(0, 5, &w, sizeof(w)));
  c->send(0, 5, &v, sizeof(v));
  v = 43;
  c->send(0, 5, &v, sizeof(v));
  REQUIRE(c->try_recv(0, 5, &w, sizeof(w)));
  REQUIRE(w == 42); // FIFO per (src, tag)
  c->recv(0, 5, &w, sizeof(w));
  REQUIRE(w == 43);
  REQUIRE(c->allreduce_max(3.5) == 3.5);
  REQUIRE(c->allreduce_sum(7) == 7);
  c->barrier();
}

TEST_CASE("tcp comm on loopback, 3 ranks", "[comm]") {
  const int n = 3;
  const int port = 23000 + int(::getpid() % 2000);
  std::vector<int> ok(static_cast<size_t>(n), 0);
  std::vector<std::thread> th;
  for (int r = 0; r < n; ++r)
    th.emplace_back([&, r] {
      std::unique_ptr<stencil::Comm> c = stencil::make_tcp_comm(r, n, "127.0.0.1", port);
      bool good = c->rank() == r && c->size() == n;
      /* allgather */
      int mine = 100 + r;
      std::vector<int> all(static_cast<size_t>(n));
      c->allgather(&mine, all.data(), sizeof(int));
      for (int i = 0; i < n; ++i) good &= all[size_t(i)] == 100 + i;
      /* allgatherv: rank r contributes r+1 ints */
      std::vector<size_t> counts;
      int total = 0;
      for (int i = 0; i < n; ++i) {
        counts.push_back(size_t(i + 1) * sizeof(int));
        total += i + 1;
      }
      std::vector<int> contrib(static_cast<size_t>(r + 1), r), gathered(static_cast<size_t>(total), -1);
      c->allgatherv(contrib.data(), gathered.data(), counts);
      good &= gathered[0] == 0 && gathered[1] == 1 && gathered[2] == 1 && gathered[5] == 2;
      /* bcast from a non-zero root */
      int b = (r == 2) ? 77 : 0;
      c->bcast(&b, sizeof(b), 2);
      good &= b == 77;
      /* ring p2p, including a relayed 1 -> 2 message, with a payload larger than a socket buffer */
      std::vector<char> big(1 << 20, char(r)), got(1 << 20, 0);
      c->send((r + 1) % n, 9, big.data(), big.size());
      c->recv((r + n - 1) % n, 9, got.data(), got.size());
      good &= got[12345] == char((r + n - 1) % n);
      good &= c->allreduce_max(double(r)) == 2.0;
      good &= c->allreduce_sum(uint64_t(r + 1)) == 6;
      c->barrier();
      ok[size_t(r)] = good;
    });
  for (auto &t : th) t.join();
  for (int r = 0; r < n; ++r) REQUIRE(ok[size_t(r)] == 1);
}

TEST_CASE("node topology from hostnames", "[comm]") {
  MpiTopology t({"a", "a", "b", "b", "a"}, 4);
  REQUIRE(t.size() == 5);
  REQUIRE(t.num_nodes() == 2);
  REQUIRE(t.colocated_size() == 3);
  REQUIRE(t.colocated_rank() == 2);
  REQUIRE(t.colocated(0));
  REQUIRE(!t.colocated(2));
  REQUIRE(t.node_of(3) == 1);
  std::unique_ptr<stencil::Comm> c = stencil::make_self_comm();
  MpiTopology self(*c);
  REQUIRE(self.colocated_size() == 1);
  REQUIRE(self.colocated_rank() == 0);
}

TEST_CASE("boundary flags", "[boundary]") {
  Boundary b;
  REQUIRE(b == Boundary::periodic());
  REQUIRE(b.dir(1, 0, 0));
  b.set_dir(1, 0, 0, false);
  b.dir(0, -1, 0) = false;
  REQUIRE(!b.dir(1, 0, 0));
  REQUIRE(!b.dir(0, -1, 0));
  REQUIRE(b.dir(-1, 0, 0));
  REQUIRE(!(b == Boundary::periodic()));
  REQUIRE(!Boundary::none().dir(1, 1, 1));
}

#include <atomic>
#include <chrono>
#include <thread>
#include "stencil/detail/launch_pool.hpp"

TEST_CASE("launch pool runs every index exactly once per call, spinning or asleep", "[misc][pool]") {
  stencil::detail::LaunchPool pool;
  std::vector<std::atomic<int>> hits(6);
  for (auto &h : hits) h = 0;
  std::atomic<int> concurrent{0}, peak{0};
  const std::function<void(size_t)> fn = [&](size_t i) {
    const int now = ++concurrent;
    int p = peak.load();
    while (now > p && !peak.compare_exchange_weak(p, now)) {
    }
    ++hits[i];
    std::this_thread::sleep_for(std::chrono::microseconds(200));
    --concurrent;
  };
  for (int round = 0; round < 50; ++round) pool.run(6, fn); // back to back: workers are still spinning
  std::this_thread::sleep_for(std::chrono::milliseconds(20)); // long enough for the workers to fall asleep
  for (int round = 0; round < 5; ++round) {
    pool.run(6, fn);
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
  pool.run(3, fn); // fewer indices than workers
  pool.run(1, fn);
  for (size_t i = 0; i < 6; ++i) REQUIRE(hits[i] == 55 + (i < 3 ? 1 : 0) + (i < 1 ? 1 : 0));
  REQUIRE(peak.load() > 1); // they did overlap
}

