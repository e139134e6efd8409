/* bench-exchange: exchange()+swap() time for five radius shapes on a domain of the given WHOLE size.
   CLI and table follow the reference's bin/bench_exchange.cu:82-98,57-64,125-195:
     --iters (30) --x --y --z (128) --q quantities (1) --fr --er --cr (2)
     name,count,trimean (S),trimean (B/s),stddev,min,avg,max
     <x>-<y>-<z>/px/<fR> | /x/<fR> | /faces/<fR> | /face&edge/<fR>/<eR> | /uniform/<fR>
   Differences: every method incl. NvlinkDirect is enabled unless method flags are given; the max-over-ranks time is
   what gets recorded (the reference records rank 0's own time). */
#include <iostream>
#include <sstream>

#include "common.hpp"
#include "stencil/statistics.hpp"

static std::pair<Statistics, uint64_t> bench(stencil::Comm &comm, size_t nIters, size_t nQuants, const Dim3 &ext, const Radius &radius,
                                             MethodFlags methods) {
  DistributedDomain dd(size_t(ext.x), size_t(ext.y), size_t(ext.z));
  dd.set_radius(radius);
  dd.set_plan_file(false);
  for (size_t i = 0; i < nQuants; ++i) dd.add_data<float>("d");
  dd.set_methods(methods);
  dd.realize();
  Statistics stats;
  for (size_t i = 0; i < nIters; ++i) {
    comm.barrier();
    const double start = stencil::Comm::wtime();
    dd.exchange();
    dd.swap();
    stats.insert(comm.allreduce_max(stencil::Comm::wtime() - start));
  }
  return std::make_pair(stats, dd.exchange_bytes_for_method(MethodFlags::All));
}

static void report(const std::string &cfg, uint64_t bytes, const Statistics &s) {
  std::cout << std::scientific << cfg << "," << s.count() << "," << s.trimean() << "," << double(bytes) / s.trimean() << "," << s.stddev() << ","
            << s.min() << "," << s.avg() << "," << s.max() << "\n"
            << std::defaultfloat;
}

int main(int argc, char **argv) {
  stencil::Comm &comm = stencil::world();
  const int rank = comm.rank();
  size_t nIters = 30, nQuants = 1, x = 128, y = 128, z = 128, fR = 2, eR = 2, cR = 2;
  MethodArgs margs;
  argparse::Parser p("halo exchange time by radius shape");
  p.add_option(nIters, "--iters");
  p.add_option(nQuants, "--q");
  p.add_option(x, "--x");
  p.add_option(y, "--y");
  p.add_option(z, "--z");
  p.add_option(fR, "--fr");
  p.add_option(eR, "--er");
  p.add_option(cR, "--cr");
  margs.add_to(p);
  if (!p.parse(argc, argv)) {
    if (rank == 0) std::cout << p.help();
    return EXIT_FAILURE;
  }
  if (p.need_help()) {
    if (rank == 0) std::cout << p.help();
    return 0;
  }
  (void)cR;
  const Dim3 ext(static_cast<int64_t>(x), static_cast<int64_t>(y), static_cast<int64_t>(z));
  const MethodFlags methods = margs.flags();
  std::ostringstream base;
  base << ext.x << "-" << ext.y << "-" << ext.z;
  if (rank == 0) std::cout << "name,count,trimean (S),trimean (B/s),stddev,min,avg,max\n";

  struct Case {
    std::string name;
    Radius radius;
  };
  std::vector<Case> cases;
  Radius r = Radius::constant(0);
  r.dir(1, 0, 0) = fR;
  cases.push_back({base.str() + "/px/" + std::to_string(fR), r});
  r.dir(-1, 0, 0) = fR;
  cases.push_back({base.str() + "/x/" + std::to_string(fR), r});
  r = Radius::constant(0);
  r.set_face(fR);
  cases.push_back({base.str() + "/faces/" + std::to_string(fR), r});
  r = Radius::constant(fR); // as the reference: faces AND edges at fR, corners at eR
  r.set_corner(eR);
  cases.push_back({base.str() + "/face&edge/" + std::to_string(fR) + "/" + std::to_string(eR), r});
  cases.push_back({base.str() + "/uniform/" + std::to_string(fR), Radius::constant(2)});
  for (const Case &c : cases) {
    auto res = bench(comm, nIters, nQuants, ext, c.radius, methods);
    if (rank == 0) report(c.name, res.second, res.first);
  }
  stencil::finalize_world();
  return 0;
}
