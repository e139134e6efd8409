/* MpiComm (src/mpi_comm.cpp) is only compiled into the library with -DUSE_MPI=ON, and this image has no MPI. To
   keep its logic exercised everywhere, the translation unit is compiled here against the single-rank MPI stand-in
   the reference arm is built with (baseline/mpi_shim/mpi.h): eager self-sends, probing receives, size checks,
   tag folding and the collectives of a one-rank world. */
#define STENCIL_USE_MPI 1
#define make_mpi_comm make_mpi_comm_over_shim
#include "../src/mpi_comm.cpp"
#undef make_mpi_comm

#include "harness.hpp"
#include <numeric>

TEST_CASE("MpiComm over the one-rank shim: p2p", "[comm][mpi]") {
  std::unique_ptr<stencil::Comm> c = stencil::make_mpi_comm_over_shim();
  REQUIRE(c != nullptr);
  REQUIRE(c->rank() == 0);
  REQUIRE(c->size() == 1);
  REQUIRE(std::string(c->name()) == "mpi");
  REQUIRE(!c->device_buffers());

  char in[64], out[64];
  REQUIRE(!c->try_recv(0, 5, out, 64));
  /* eager: two sends on one tag complete without a receiver and arrive in order; other tags do not interfere */
  for (int i = 0; i < 64; ++i) in[i] = char(i);
  c->send(0, 5, in, 64);
  for (int i = 0; i < 64; ++i) in[i] = char(100 + i);
  c->send(0, 5, in, 64);
  const int other = 77;
  c->send(0, 6, &other, sizeof(other));
  REQUIRE(c->try_recv(0, 5, out, 64));
  REQUIRE(out[0] == 0);
  REQUIRE(out[63] == 63);
  c->recv(0, 5, out, 64);
  REQUIRE(out[0] == 100);
  int got = 0;
  REQUIRE(c->try_recv(0, 6, &got, sizeof(got)));
  REQUIRE(got == 77);
  REQUIRE(!c->try_recv(0, 6, &got, sizeof(got)));

  /* a tag above the guaranteed MPI_TAG_UB still round-trips (both sides fold it the same way) */
  const int big = (3 << 20) | 12345;
  c->send(0, big, &other, sizeof(other));
  got = 0;
  REQUIRE(c->try_recv(0, big, &got, sizeof(got)));
  REQUIRE(got == 77);
}

TEST_CASE("MpiComm over the one-rank shim: collectives", "[comm][mpi]") {
  std::unique_ptr<stencil::Comm> c = stencil::make_mpi_comm_over_shim();
  c->barrier();
  int v = 42, all = 0;
  c->allgather(&v, &all, sizeof(int));
  REQUIRE(all == 42);
  std::vector<int> mine = {1, 2, 3}, gathered(3, 0);
  c->allgatherv(mine.data(), gathered.data(), {3 * sizeof(int)});
  REQUIRE(std::accumulate(gathered.begin(), gathered.end(), 0) == 6);
  c->bcast(&v, sizeof(int), 0);
  REQUIRE(v == 42);
  REQUIRE(c->allreduce_max(2.5) == 2.5);
  REQUIRE(c->allreduce_sum(uint64_t(9)) == 9u);
}
