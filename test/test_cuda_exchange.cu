/* Full DistributedDomain exchanges, checked on the ENTIRE allocation (halos included) against the periodic
   analytic field. Ground covered by the reference's test_exchange.cu / test_cuda_mpi_exchange.cu /
   test_cuda_mpi_distributed_domain.cu (radius 0,1,2, +x=2, -x=1, both; 2 subdomains on GPU 0; staged transport
   forced for every message incl. self-sends) and the gaps SURVEY section 4 lists: every method, several quantities
   of mixed size, uneven partitions, interior/exterior, paraview, byte counters, plan file. */
#include "harness.hpp"
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <vector>
#include "stencil/stencil.hpp"

/* MpiComm compiled against the one-rank MPI stand-in, as in test_cpu_mpi_comm.cpp */
#define STENCIL_USE_MPI 1
#define make_mpi_comm make_mpi_comm_over_shim_cuda
#include "../src/mpi_comm.cpp"
#undef make_mpi_comm

namespace {

/* a one-rank communicator that takes DEVICE pointers, standing in for a CUDA-aware MPI */
class DeviceLoopComm : public stencil::Comm {
public:
  int rank() const override { return 0; }
  int size() const override { return 1; }
  const char *name() const override { return "device-loop"; }
  bool device_buffers() const override { return true; }
  void send(int, int tag, const void *buf, size_t bytes) override {
    std::vector<char> m(bytes);
    CUDA_RUNTIME(cudaMemcpy(m.data(), buf, bytes, cudaMemcpyDefault));
    box_[tag].push_back(std::move(m));
    ++sends;
  }
  bool try_recv(int, int tag, void *buf, size_t bytes) override {
    auto it = box_.find(tag);
    if (it == box_.end() || it->second.empty()) return false;
    if (it->second.front().size() != bytes) return false;
    CUDA_RUNTIME(cudaMemcpy(buf, it->second.front().data(), bytes, cudaMemcpyDefault));
    it->second.erase(it->second.begin());
    return true;
  }
  size_t sends = 0;

private:
  std::map<int, std::vector<std::vector<char>>> box_;
};

/* exact, quantity-specific encodings of a global coordinate */
float enc_f(const Dim3 &p) { return float(p.x + 64 * p.y + 4096 * p.z); }
double enc_d(const Dim3 &p) { return double(p.x) + 1e3 * double(p.y) + 1e6 * double(p.z) + 0.25; }
unsigned char enc_c(const Dim3 &p) { return (unsigned char)((p.x * 7 + p.y * 13 + p.z * 29) & 0xFF); }

struct Fields {
  DataHandle<float> f;
  DataHandle<unsigned char> c;
  DataHandle<double> d;
};

void fill(DistributedDomain &dd) {
  for (LocalDomain &ld : dd.domains()) {
    const Dim3 raw = ld.raw_size(), sz = ld.size(), org = ld.origin();
    /* poison the whole allocation, then write the compute region */
    std::vector<float> pf(size_t(raw.flatten()), -1.f);
    std::vector<unsigned char> pc(size_t(raw.flatten()), 0xEE);
    std::vector<double> pd(size_t(raw.flatten()), -1.0);
    ld.region_from_host(pf.data(), Dim3(0, 0, 0), raw, 0);
    ld.region_from_host(pc.data(), Dim3(0, 0, 0), raw, 1);
    ld.region_from_host(pd.data(), Dim3(0, 0, 0), raw, 2);
    std::vector<float> vf(size_t(sz.flatten()));
    std::vector<unsigned char> vc(size_t(sz.flatten()));
    std::vector<double> vd(size_t(sz.flatten()));
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x, ++i) {
          const Dim3 p = org + Dim3(x, y, z);
          vf[i] = enc_f(p);
          vc[i] = enc_c(p);
          vd[i] = enc_d(p);
        }
    const Dim3 pos = ld.halo_pos(Dim3(0, 0, 0), true);
    ld.region_from_host(vf.data(), pos, sz, 0);
    ld.region_from_host(vc.data(), pos, sz, 1);
    ld.region_from_host(vd.data(), pos, sz, 2);
  }
}

/* every cell of every allocation: compute cells keep their value; halo cells on side d hold the periodic image iff
   radius.dir(d) > 0, else they must still be poisoned */
size_t count_wrong(DistributedDomain &dd, const Radius &radius) {
  size_t wrong = 0;
  const Dim3 global = dd.size();
  for (LocalDomain &ld : dd.domains()) {
    const Dim3 raw = ld.raw_size(), sz = ld.size(), horg = ld.halo_origin();
    const std::vector<unsigned char> bf = ld.quantity_to_host(0), bc = ld.quantity_to_host(1), bd = ld.quantity_to_host(2);
    const float *vf = reinterpret_cast<const float *>(bf.data());
    const double *vd = reinterpret_cast<const double *>(bd.data());
    const int64_t lo[3] = {int64_t(radius.x(-1)), int64_t(radius.y(-1)), int64_t(radius.z(-1))};
    size_t i = 0;
    for (int64_t z = 0; z < raw.z; ++z)
      for (int64_t y = 0; y < raw.y; ++y)
        for (int64_t x = 0; x < raw.x; ++x, ++i) {
          const int64_t c[3] = {x, y, z};
          Dim3 side;
          for (int a = 0; a < 3; ++a) side[a] = c[a] < lo[a] ? -1 : (c[a] >= lo[a] + sz[a] ? 1 : 0);
          const Dim3 p = (horg + Dim3(x, y, z)).wrapped(global);
          const bool expectData = side == Dim3(0, 0, 0) || radius.dir(side) > 0;
          if (expectData) {
            wrong += vf[i] != enc_f(p);
            wrong += bc[i] != enc_c(p);
            wrong += vd[i] != enc_d(p);
          } else {
            wrong += vf[i] != -1.f;
            wrong += bc[i] != 0xEE;
            wrong += vd[i] != -1.0;
          }
        }
  }
  return wrong;
}

size_t run_case(const Dim3 &size, const Radius &radius, MethodFlags methods, const std::vector<int> &gpus, PlacementStrategy strategy,
                int rounds = 2, stencil::Comm *comm = nullptr, uint64_t *awareBytes = nullptr, uint64_t *stagedBytes = nullptr) {
  DistributedDomain dd(size_t(size.x), size_t(size.y), size_t(size.z), comm ? *comm : stencil::world());
  dd.set_gpus(gpus);
  dd.set_radius(radius);
  dd.set_methods(methods);
  dd.set_placement(strategy);
  dd.set_plan_file(false);
  Fields q{dd.add_data<float>("f"), dd.add_data<unsigned char>("c"), dd.add_data<double>("d")};
  (void)q;
  dd.realize();
  if (awareBytes) *awareBytes = dd.exchange_bytes_for_method(MethodFlags::CudaAwareMpi);
  if (stagedBytes) *stagedBytes = dd.exchange_bytes_for_method(MethodFlags::CudaMpi);
  size_t wrong = 0;
  for (int r = 0; r < rounds; ++r) {
    fill(dd);
    dd.exchange();
    wrong += count_wrong(dd, radius);
    dd.swap(); // next round runs on the other buffers
  }
  return wrong;
}

std::vector<Radius> radius_shapes() {
  std::vector<Radius> v;
  v.push_back(Radius::constant(1));
  v.push_back(Radius::constant(2));
  Radius faces = Radius::constant(0);
  faces.set_face(1);
  v.push_back(faces);
  Radius px = Radius::constant(0);
  px.dir(1, 0, 0) = 2;
  v.push_back(px);
  Radius mx = Radius::constant(0);
  mx.dir(-1, 0, 0) = 1;
  v.push_back(mx);
  Radius both = Radius::constant(0);
  both.dir(1, 0, 0) = 2;
  both.dir(-1, 0, 0) = 1;
  v.push_back(both);
  v.push_back(Radius::face_edge_corner(3, 2, 1));
  Radius lop = Radius::constant(1); // uneven on every axis, edges and corners on
  lop.dir(1, 0, 0) = 3;
  lop.dir(0, -1, 0) = 2;
  lop.dir(0, 0, 1) = 2;
  v.push_back(lop);
  return v;
}

/* tests written against a grid that cuts x ask for the reference's partition order */
struct CutX {
  stencil::detail::TieBreak saved = partition_tie_break();
  CutX() { set_partition_tie_break(stencil::detail::TieBreak::X); }
  ~CutX() { set_partition_tie_break(saved); }
};

int device_count() {
  int n = 0;
  cudaGetDeviceCount(&n);
  return n;
}

} // namespace

TEST_CASE("exchange: radius 0 is a no-op", "[cuda][exchange]") {
  REQUIRE(run_case(Dim3(10, 10, 10), Radius::constant(0), MethodFlags::All, {0, 0}, PlacementStrategy::NodeAware) == 0);
}

TEST_CASE("exchange: one subdomain wraps onto itself, all methods", "[cuda][exchange]") {
  for (const Radius &r : radius_shapes()) {
    REQUIRE(run_case(Dim3(10, 10, 10), r, MethodFlags::All, {0}, PlacementStrategy::NodeAware) == 0);
    REQUIRE(run_case(Dim3(10, 10, 10), r, MethodFlags::CudaMpi, {0}, PlacementStrategy::Trivial) == 0);
    REQUIRE(run_case(Dim3(10, 10, 10), r, MethodFlags::NvlinkDirect, {0}, PlacementStrategy::Trivial) == 0);
  }
}

TEST_CASE("exchange: two subdomains on GPU 0, every method", "[cuda][exchange]") {
  const MethodFlags methods[] = {MethodFlags::All,
                                 MethodFlags::CudaMpi,
                                 MethodFlags::CudaKernel | MethodFlags::CudaMpi,
                                 MethodFlags::CudaMemcpyPeer,
                                 MethodFlags::NvlinkDirect,
                                 MethodFlags::CudaKernel};
  for (MethodFlags m : methods)
    for (const Radius &r : radius_shapes()) REQUIRE(run_case(Dim3(10, 10, 10), r, m, {0, 0}, PlacementStrategy::NodeAware) == 0);
}

TEST_CASE("exchange: uneven partitions, 3/4/6/8 subdomains on GPU 0", "[cuda][exchange]") {
  const Radius r = Radius::face_edge_corner(2, 2, 1);
  REQUIRE(run_case(Dim3(11, 7, 9), r, MethodFlags::All, {0, 0, 0}, PlacementStrategy::NodeAware) == 0);
  REQUIRE(run_case(Dim3(13, 9, 7), r, MethodFlags::All, {0, 0, 0, 0}, PlacementStrategy::Trivial) == 0);
  REQUIRE(run_case(Dim3(17, 11, 9), r, MethodFlags::CudaMpi, {0, 0, 0, 0, 0, 0}, PlacementStrategy::NodeAware) == 0);
  REQUIRE(run_case(Dim3(15, 14, 13), Radius::constant(3), MethodFlags::All, {0, 0, 0, 0, 0, 0, 0, 0}, PlacementStrategy::NodeAware) == 0);
  REQUIRE(run_case(Dim3(15, 14, 13), Radius::constant(3), MethodFlags::CudaMemcpyPeer, {0, 0, 0, 0, 0, 0, 0, 0}, PlacementStrategy::NodeAware) == 0);
}

TEST_CASE("exchange: bigger box exercises vector paths and many chunks", "[cuda][exchange]") {
  REQUIRE(run_case(Dim3(128, 96, 64), Radius::constant(3), MethodFlags::All, {0, 0}, PlacementStrategy::NodeAware, 1) == 0);
  REQUIRE(run_case(Dim3(130, 66, 34), Radius::constant(2), MethodFlags::NvlinkDirect, {0, 0, 0, 0}, PlacementStrategy::NodeAware, 1) == 0);
}

TEST_CASE("exchange: all visible GPUs, direct NVLink stores and the legacy peer path", "[cuda][exchange][multigpu]") {
  const int n = device_count();
  if (n < 2) {
    std::printf("  (1 GPU visible: multi-GPU exchange not exercised)\n");
    return;
  }
  std::vector<int> gpus;
  for (int i = 0; i < n; ++i) gpus.push_back(i);
  {
    /* the reference's partition cuts x across GPUs: thin strided faces over NVLink (mailboxes + scatter) */
    const stencil::detail::TieBreak saved = partition_tie_break();
    set_partition_tie_break(stencil::detail::TieBreak::X);
    size_t wrong = 0;
    for (const Radius &r : radius_shapes()) wrong += run_case(Dim3(40, 36, 28), r, MethodFlags::All, gpus, PlacementStrategy::NodeAware);
    wrong += run_case(Dim3(96, 20, 12), Radius::constant(3), MethodFlags::All, gpus, PlacementStrategy::NodeAware);
    set_partition_tie_break(saved);
    REQUIRE(wrong == 0);
  }
  for (const Radius &r : radius_shapes()) {
    REQUIRE(run_case(Dim3(40, 36, 28), r, MethodFlags::All, gpus, PlacementStrategy::NodeAware) == 0);
    REQUIRE(run_case(Dim3(40, 36, 28), r, MethodFlags::CudaMemcpyPeer | MethodFlags::CudaKernel, gpus, PlacementStrategy::NodeAware) == 0);
    REQUIRE(run_case(Dim3(40, 36, 28), r, MethodFlags::CudaMpi, gpus, PlacementStrategy::Trivial) == 0);
  }
  /* two subdomains per GPU */
  std::vector<int> twice = gpus;
  twice.insert(twice.end(), gpus.begin(), gpus.end());
  REQUIRE(run_case(Dim3(40, 36, 28), Radius::constant(2), MethodFlags::All, twice, PlacementStrategy::NodeAware) == 0);
}

TEST_CASE("interior and exterior regions tile the compute region", "[cuda][exchange]") {
  DistributedDomain dd(20, 18, 16);
  dd.set_gpus({0, 0});
  Radius r = Radius::constant(0);
  r.set_face(1);
  r.dir(1, 0, 0) = 3;
  r.dir(0, -1, 0) = 2;
  dd.set_radius(r);
  dd.set_plan_file(false);
  dd.add_data<float>("f");
  dd.realize();
  const std::vector<Rect3> in = dd.get_interior();
  const std::vector<std::vector<Rect3>> ex = dd.get_exterior();
  REQUIRE(in.size() == 2);
  for (size_t di = 0; di < 2; ++di) {
    const Rect3 com = dd.domains()[di].get_compute_region();
    REQUIRE(in[di].lo == com.lo + Dim3(1, 2, 1));
    REQUIRE(in[di].hi == com.hi - Dim3(3, 1, 1));
    REQUIRE(ex[di].size() == 6);
    /* +x slab first and spanning the whole y/z range */
    REQUIRE(ex[di][0] == Rect3(Dim3(in[di].hi.x, com.lo.y, com.lo.z), com.hi));
    int64_t cells = in[di].extent().flatten();
    for (const Rect3 &s : ex[di]) cells += s.extent().flatten();
    REQUIRE(cells == com.extent().flatten());
    /* pairwise disjoint */
    std::vector<Rect3> all = ex[di];
    all.push_back(in[di]);
    for (size_t a = 0; a < all.size(); ++a)
      for (size_t b = a + 1; b < all.size(); ++b) {
        bool overlap = true;
        for (int k = 0; k < 3; ++k) overlap &= all[a].lo[k] < all[b].hi[k] && all[b].lo[k] < all[a].hi[k];
        REQUIRE(!overlap);
      }
  }
  REQUIRE(dd.get_compute_region() == Rect3(Dim3(0, 0, 0), Dim3(20, 18, 16)));
}

TEST_CASE("byte counters, plan file and paraview output", "[cuda][exchange]") {
  CutX grid;
  DistributedDomain dd(8, 6, 4);
  dd.set_gpus({0, 0});
  dd.set_radius(1);
  dd.set_methods(MethodFlags::CudaKernel | MethodFlags::CudaMpi);
  auto hf = dd.add_data<float>("temp");
  dd.add_data<double>();
  dd.realize();
  /* 2x1x1 grid of 4x6x4: every one of the 26 messages of each subdomain stays on GPU 0 => all "kernel" */
  uint64_t perDomain = 0;
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; ++x)
        if (x || y || z) perDomain += uint64_t((x ? 1 : 4) * (y ? 1 : 6) * (z ? 1 : 4)) * 12;
  REQUIRE(dd.exchange_bytes_for_method(MethodFlags::CudaKernel) == 2 * perDomain);
  REQUIRE(dd.exchange_bytes_for_method(MethodFlags::CudaMpi) == 0);
  REQUIRE(dd.exchange_bytes_for_method(MethodFlags::All) == 2 * perDomain);
  std::ifstream plan("plan_0.txt");
  REQUIRE(plan.good());
  std::stringstream ss;
  ss << plan.rdbuf();
  REQUIRE(ss.str().find("rank=0") != std::string::npos);
  REQUIRE(ss.str().find("== peerAccess ==") != std::string::npos);
  REQUIRE(ss.str().find("0:cuda0:[0,0,0] sz=[4,6,4]") != std::string::npos);

  for (LocalDomain &ld : dd.domains()) {
    const Dim3 sz = ld.size();
    std::vector<float> v(size_t(sz.flatten()));
    std::vector<double> w(size_t(sz.flatten()), 2.5);
    size_t i = 0;
    for (int64_t z = 0; z < sz.z; ++z)
      for (int64_t y = 0; y < sz.y; ++y)
        for (int64_t x = 0; x < sz.x; ++x) v[i++] = enc_f(ld.origin() + Dim3(x, y, z));
    ld.region_from_host(v.data(), ld.halo_pos(Dim3(0, 0, 0), true), sz, 0);
    ld.region_from_host(w.data(), ld.halo_pos(Dim3(0, 0, 0), true), sz, 1);
  }
  (void)hf;
  dd.write_paraview("/tmp/stencil_test_pv");
  std::ifstream pv("/tmp/stencil_test_pv_1.txt");
  REQUIRE(pv.good());
  std::string header, first;
  std::getline(pv, header);
  std::getline(pv, first);
  REQUIRE(header == "Z,Y,X,temp,data1");
  REQUIRE(first == "0,0,4,4.000000,2.500000"); // subdomain 1 starts at x=4
  size_t lines = 2;
  std::string l;
  while (std::getline(pv, l)) ++lines;
  REQUIRE(lines == 1 + 4 * 6 * 4);
}

TEST_CASE("non-periodic boundaries leave the outer halos untouched", "[cuda][exchange]") {
  CutX grid;
  DistributedDomain dd(12, 6, 6);
  dd.set_gpus({0, 0});
  dd.set_radius(1);
  dd.set_plan_file(false);
  Boundary b; // periodic in y and z only
  b.dir(1, 0, 0) = false;
  b.dir(-1, 0, 0) = false;
  dd.set_boundary(b);
  dd.add_data<float>("f");
  dd.add_data<unsigned char>("c");
  dd.add_data<double>("d");
  dd.realize();
  fill(dd);
  dd.exchange();
  for (LocalDomain &ld : dd.domains()) {
    const std::vector<unsigned char> bytes = ld.quantity_to_host(0);
    const float *v = reinterpret_cast<const float *>(bytes.data());
    const Dim3 raw = ld.raw_size();
    const bool leftmost = ld.origin().x == 0;
    for (int64_t z = 1; z < raw.z - 1; ++z)
      for (int64_t y = 1; y < raw.y - 1; ++y) {
        const float outer = v[size_t((z * raw.y + y) * raw.x + (leftmost ? 0 : raw.x - 1))];
        const float inner = v[size_t((z * raw.y + y) * raw.x + (leftmost ? raw.x - 1 : 0))];
        REQUIRE(outer == -1.f); // the global x boundary: nothing arrives
        REQUIRE(inner != -1.f); // the internal interface is exchanged
      }
  }
}

TEST_CASE("exchange: staged transport over MpiComm, and device-pointer messages (CudaAwareMpi)", "[cuda][exchange][comm]") {
  /* the reference forces every message, self-sends included, through MPI in test_cuda_mpi_exchange.cu */
  std::unique_ptr<stencil::Comm> mpi = stencil::make_mpi_comm_over_shim_cuda();
  uint64_t aware = 1, staged = 0;
  REQUIRE(run_case(Dim3(12, 10, 9), Radius::constant(2), MethodFlags::CudaMpi, {0, 0}, PlacementStrategy::Trivial, 3, mpi.get(), &aware, &staged) == 0);
  REQUIRE(aware == 0);
  REQUIRE(staged > 0);
  /* asking for CudaAwareMpi on a communicator without device pointers falls back to pinned staging */
  REQUIRE(run_case(Dim3(12, 10, 9), Radius::constant(1), MethodFlags::CudaAwareMpi, {0, 0}, PlacementStrategy::Trivial, 2, mpi.get(), &aware, &staged) == 0);
  REQUIRE(aware == 0);
  REQUIRE(staged > 0);

  DeviceLoopComm dev;
  for (const Radius &r : radius_shapes()) {
    REQUIRE(run_case(Dim3(10, 10, 10), r, MethodFlags::CudaAwareMpi, {0, 0}, PlacementStrategy::Trivial, 2, &dev, &aware, &staged) == 0);
    REQUIRE(staged == 0);
    REQUIRE(aware > 0);
  }
  REQUIRE(dev.sends > 0);
  /* with faster methods enabled too, only what they cannot carry is left to the communicator (nothing, on one node) */
  const size_t before = dev.sends;
  REQUIRE(run_case(Dim3(10, 10, 10), Radius::constant(1), MethodFlags::All | MethodFlags::CudaAwareMpi, {0, 0}, PlacementStrategy::Trivial, 2, &dev, &aware, &staged) == 0);
  REQUIRE(dev.sends == before);
  REQUIRE(aware == 0);
}

TEST_CASE("exchange_async / exchange_wait and the event-chained form", "[cuda][exchange][async]") {
  const Radius radius = Radius::constant(2);
  for (MethodFlags methods : {MethodFlags::All, MethodFlags::CudaMpi, MethodFlags::CudaMemcpyPeer | MethodFlags::CudaMpi}) {
    DistributedDomain dd(14, 12, 10);
    dd.set_gpus({0, 0, 0});
    dd.set_radius(radius);
    dd.set_methods(methods);
    dd.set_plan_file(false);
    Fields q{dd.add_data<float>("f"), dd.add_data<unsigned char>("c"), dd.add_data<double>("d")};
    (void)q;
    dd.realize();
    /* staged messages need the host; direct stores and peer copies do not: device-only is a property of the plan */
    REQUIRE(dd.exchange_device_only() == (methods != MethodFlags::CudaMpi));
    for (int round = 0; round < 3; ++round) {
      fill(dd);
      dd.exchange_async();
      dd.exchange_wait();
      REQUIRE(count_wrong(dd, radius) == 0);
      dd.swap();
    }
    dd.exchange_wait(); // nothing in flight: no-op
  }

  /* never waiting on the host for the exchange: user streams are chained to it by events in both directions */
  DistributedDomain dd(16, 12, 12);
  dd.set_gpus({0, 0});
  dd.set_radius(radius);
  dd.set_plan_file(false);
  Fields q{dd.add_data<float>("f"), dd.add_data<unsigned char>("c"), dd.add_data<double>("d")};
  (void)q;
  dd.realize();
  REQUIRE(dd.exchange_device_only());
  const size_t nd = dd.domains().size();
  std::vector<cudaStream_t> user(nd);
  std::vector<cudaEvent_t> mine(nd);
  for (size_t di = 0; di < nd; ++di) {
    CUDA_RUNTIME(cudaSetDevice(dd.domains()[di].gpu()));
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&user[di], cudaStreamNonBlocking));
    CUDA_RUNTIME(cudaEventCreateWithFlags(&mine[di], cudaEventDisableTiming));
  }
  for (int round = 0; round < 4; ++round) {
    fill(dd); // synchronous host copies stand in for the user's kernels
    for (size_t di = 0; di < nd; ++di) {
      CUDA_RUNTIME(cudaEventRecord(mine[di], user[di]));
      dd.exchange_after(di, mine[di]);
    }
    dd.exchange_async();
    for (size_t di = 0; di < nd; ++di) CUDA_RUNTIME(cudaStreamWaitEvent(user[di], dd.exchange_done_event(di), 0));
    for (size_t di = 0; di < nd; ++di) CUDA_RUNTIME(cudaStreamSynchronize(user[di]));
    REQUIRE(count_wrong(dd, radius) == 0);
    dd.swap();
  }
  dd.exchange_wait();
  for (size_t di = 0; di < nd; ++di) {
    cudaStreamDestroy(user[di]);
    cudaEventDestroy(mine[di]);
  }
}
