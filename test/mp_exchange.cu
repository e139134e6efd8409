/* test_mp_exchange: one rank of a multi-process exchange test (the counterpart of the reference's test_cuda_mpi
   binaries: test_cuda_mpi_exchange.cu, test_cuda_mpi_distributed_domain.cu, test_cuda_mpi_colocatedtx.cu,
   test_cuda_mpi_cudaipc.cu). Start N copies with RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT set (torchrun does).
   Every rank fills its subdomains with an exact encoding of the global coordinate, exchanges with the requested
   methods, and checks its whole allocations (halos included) against the periodic image.
     test_mp_exchange <methods bitmask> [x y z] [radius] [rounds]
   Exit code 0 and a line "rank R ok" on success. Works with several ranks sharing one GPU (then the direct path is
   not eligible between them and the staged / colocated transports carry the traffic). */
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "stencil/stencil.hpp"

static float enc_f(const Dim3 &p) { return float(p.x + 64 * p.y + 4096 * p.z); }
static double enc_d(const Dim3 &p) { return double(p.x) + 1e3 * double(p.y) + 1e6 * double(p.z) + 0.5; }

int main(int argc, char **argv) {
  const int methods = argc > 1 ? std::atoi(argv[1]) : int(MethodFlags::All);
  const size_t x = argc > 4 ? size_t(std::atoi(argv[2])) : 24, y = argc > 4 ? size_t(std::atoi(argv[3])) : 18, z = argc > 4 ? size_t(std::atoi(argv[4])) : 14;
  const int r = argc > 5 ? std::atoi(argv[5]) : 2;
  const int rounds = argc > 6 ? std::atoi(argv[6]) : 3;
  stencil::Comm &comm = stencil::world();
  size_t wrong = 0;
  {
    DistributedDomain dd(x, y, z);
    Radius radius = Radius::face_edge_corner(size_t(r), size_t(r), size_t(r > 1 ? r - 1 : r));
    dd.set_radius(radius);
    dd.set_methods(static_cast<MethodFlags>(methods));
    dd.set_plan_file(false);
    dd.add_data<float>("f");
    dd.add_data<double>("d");
    dd.realize();
    const Dim3 global = dd.size();
    for (int round = 0; round < rounds; ++round) {
      for (LocalDomain &ld : dd.domains()) {
        const Dim3 raw = ld.raw_size(), sz = ld.size(), org = ld.origin();
        std::vector<float> pf(size_t(raw.flatten()), -1.f), vf(size_t(sz.flatten()));
        std::vector<double> pd(size_t(raw.flatten()), -1.0), vd(size_t(sz.flatten()));
        size_t i = 0;
        for (int64_t c = 0; c < sz.z; ++c)
          for (int64_t b = 0; b < sz.y; ++b)
            for (int64_t a = 0; a < sz.x; ++a, ++i) {
              vf[i] = enc_f(org + Dim3(a, b, c)) + float(round);
              vd[i] = enc_d(org + Dim3(a, b, c)) + double(round);
            }
        ld.region_from_host(pf.data(), Dim3(0, 0, 0), raw, 0);
        ld.region_from_host(pd.data(), Dim3(0, 0, 0), raw, 1);
        ld.region_from_host(vf.data(), ld.halo_pos(Dim3(0, 0, 0), true), sz, 0);
        ld.region_from_host(vd.data(), ld.halo_pos(Dim3(0, 0, 0), true), sz, 1);
      }
      comm.barrier(); // nobody pushes into a subdomain that is still being initialised
      dd.exchange();
      for (LocalDomain &ld : dd.domains()) {
        const Dim3 raw = ld.raw_size(), horg = ld.halo_origin();
        const std::vector<unsigned char> bf = ld.quantity_to_host(0), bd = ld.quantity_to_host(1);
        const float *vf = reinterpret_cast<const float *>(bf.data());
        const double *vd = reinterpret_cast<const double *>(bd.data());
        size_t i = 0;
        for (int64_t c = 0; c < raw.z; ++c)
          for (int64_t b = 0; b < raw.y; ++b)
            for (int64_t a = 0; a < raw.x; ++a, ++i) {
              const Dim3 p = (horg + Dim3(a, b, c)).wrapped(global);
              wrong += vf[i] != enc_f(p) + float(round);
              wrong += vd[i] != enc_d(p) + double(round);
            }
      }
      comm.barrier();
      dd.swap();
    }
  }
  const uint64_t total = comm.allreduce_sum(uint64_t(wrong));
  std::printf("rank %d %s (%zu wrong here, %llu in the job)\n", comm.rank(), total == 0 ? "ok" : "FAILED", wrong, (unsigned long long)total);
  stencil::finalize_world();
  return total == 0 ? 0 : 1;
}
