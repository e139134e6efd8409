/* Individually constructed transports (include/stencil/tx_cuda.cuh): a ring of RemoteSender / RemoteRecver pairs between
   hand-made LocalDomains, driven through the StatefulSender / StatefulRecver phases, the way the reference's
   test/test_cuda_mpi_colocatedtx.cu wires its senders and recvers by hand. One process: messages travel through the
   self communicator (pack -> pinned host -> Comm -> pinned host -> unpack). */
#include "harness.hpp"
#include <cstring>
#include <memory>
#include <vector>
#include "stencil/tx_cuda.cuh"

namespace {
float cell_value(int dom, int64_t x, int64_t y, int64_t z) { return float(dom * 1000000 + z * 10000 + y * 100 + x); }

void fill_interior(LocalDomain &ld, int dom) {
  const Dim3 sz = ld.size(), lo = ld.halo_pos(Dim3(0, 0, 0), true), raw = ld.raw_size();
  std::vector<float> host(size_t(raw.flatten()), -1.f);
  for (int64_t z = 0; z < sz.z; ++z)
    for (int64_t y = 0; y < sz.y; ++y)
      for (int64_t x = 0; x < sz.x; ++x) host[size_t(((z + lo.z) * raw.y + y + lo.y) * raw.x + x + lo.x)] = cell_value(dom, x, y, z);
  ld.region_from_host(host.data(), Dim3(0, 0, 0), raw, 0);
}
} // namespace

TEST_CASE("tx: a ring of hand-wired RemoteSender / RemoteRecver pairs delivers +x and -x faces", "[cuda][tx]") {
  const int n = 3;
  const Dim3 sz(12, 7, 5);
  std::vector<std::unique_ptr<LocalDomain>> doms;
  for (int i = 0; i < n; ++i) {
    doms.emplace_back(new LocalDomain(sz, Dim3(i * sz.x, 0, 0), 0));
    Radius r = Radius::constant(0);
    r.dir(1, 0, 0) = 2;
    r.dir(-1, 0, 0) = 1;
    doms.back()->set_radius(r);
    doms.back()->add_data<float>();
    doms.back()->realize();
    fill_interior(*doms.back(), i);
  }
  stencil::Comm &comm = stencil::world();
  std::vector<std::unique_ptr<RemoteSender>> senders;
  std::vector<std::unique_ptr<RemoteRecver>> recvers;
  for (int i = 0; i < n; ++i)
    for (int d : {1, -1}) {
      const int dst = (i + d + n) % n;
      senders.emplace_back(new RemoteSender(comm.rank(), i, comm.rank(), dst, *doms[size_t(i)], comm));
      senders.back()->start_prepare(doms[size_t(i)].get(), {Message(Dim3(d, 0, 0), i, dst)});
      senders.back()->finish_prepare();
      recvers.emplace_back(new RemoteRecver(comm.rank(), i, comm.rank(), dst, *doms[size_t(dst)], comm));
      recvers.back()->start_prepare(doms[size_t(dst)].get(), {Message(Dim3(d, 0, 0), i, dst)});
      recvers.back()->finish_prepare();
      /* a message travelling in +x is sized by the receiver's -x halo (1 cell wide here), in -x by its +x halo (2 cells) */
      REQUIRE(senders.back()->size() == (d > 0 ? 1 : 2) * sz.y * sz.z * 4);
      REQUIRE(recvers.back()->size() == senders.back()->size());
    }
  for (int round = 0; round < 2; ++round) { // twice: the objects are reusable
    for (auto &r : recvers) r->recv();
    for (auto &s : senders) s->send();
    bool busy = true;
    while (busy) {
      busy = false;
      for (auto &s : senders)
        if (s->active()) {
          busy = true;
          if (s->next_ready()) s->next();
        }
      for (auto &r : recvers)
        if (r->active()) {
          busy = true;
          if (r->next_ready()) r->next();
        }
    }
    for (auto &s : senders) s->wait();
    for (auto &r : recvers) r->wait();
  }
  size_t bad = 0;
  for (int i = 0; i < n; ++i) {
    LocalDomain &ld = *doms[size_t(i)];
    const Dim3 raw = ld.raw_size();
    const std::vector<unsigned char> bytes = ld.quantity_to_host(0);
    const float *v = reinterpret_cast<const float *>(bytes.data());
    const int left = (i - 1 + n) % n, right = (i + 1) % n;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y) {
        /* raw x = 0 is the -x halo (1 cell): the left neighbour's last cell; raw x = 13, 14 the +x halo: the right neighbour's first two */
        const float *row = v + (z * raw.y + y) * raw.x;
        bad += row[0] != cell_value(left, sz.x - 1, y, z);
        bad += row[1 + sz.x] != cell_value(right, 0, y, z);
        bad += row[2 + sz.x] != cell_value(right, 1, y, z);
        for (int64_t x = 0; x < sz.x; ++x) bad += row[1 + x] != cell_value(i, x, y, z);
      }
  }
  REQUIRE(bad == 0);
}
