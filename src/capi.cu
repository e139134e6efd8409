/* C ABI (include/stencil_c.h): thin, allocation-free translations onto the C++ objects. */
#include "stencil_c.h"

#include <cstring>
#include <memory>

#include "stencil/copy.cuh"
#include "stencil/astaroth.hpp"
#include "stencil/jacobi.hpp"
#include "stencil/pack_kernel.cuh"
#include "stencil/reduce.hpp"
#include "stencil/stencil.hpp"

#ifndef STENCIL_GIT_REV
#define STENCIL_GIT_REV "unknown"
#endif

struct stencil_domain {
  std::unique_ptr<DistributedDomain> owned;
  DistributedDomain *dd;
};
struct stencil_jacobi {
  std::unique_ptr<stencil::Jacobi3D> j;
  stencil_domain view; // borrowed view of j->domain()
};

struct stencil_astaroth {
  std::unique_ptr<stencil::AstarothSim> a;
  stencil_domain view;
};

namespace {
Dim3 d3(const int64_t v[3]) { return Dim3(v[0], v[1], v[2]); }
void put(int64_t out[3], const Dim3 &d) {
  out[0] = d.x;
  out[1] = d.y;
  out[2] = d.z;
}
Radius radius_from(const uint64_t r[27]) {
  Radius out;
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; ++x) out.dir(x, y, z) = size_t(r[DirectionMap<int>::slot(x, y, z)]);
  return out;
}
cudaStream_t st(void *s) { return static_cast<cudaStream_t>(s); }
LocalDomain *sub(const stencil_domain_t *dd, int di) {
  if (!dd || di < 0 || size_t(di) >= dd->dd->domains().size()) return nullptr;
  return &dd->dd->domains()[size_t(di)];
}
} // namespace

extern "C" {

const char *stencil_version(void) { return "stencil-b200 " STENCIL_GIT_REV " sm_100a"; }

int stencil_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n;
}

void *stencil_host_alloc_pinned(size_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
    (void)cudaGetLastError();
    return nullptr;
  }
  std::memset(p, 0, bytes);
  return p;
}
void stencil_host_free_pinned(void *p) {
  if (p) cudaFreeHost(p);
}

int stencil_comm_init_tcp(int rank, int size, const char *addr, int port) {
  if (size < 1 || rank < 0 || rank >= size || !addr) return -1;
  if (size == 1)
    stencil::set_world(stencil::make_self_comm());
  else
    stencil::set_world(stencil::make_tcp_comm(rank, size, addr, port));
  return 0;
}
int stencil_comm_rank(void) { return stencil::world().rank(); }
int stencil_comm_size(void) { return stencil::world().size(); }
void stencil_comm_barrier(void) { stencil::world().barrier(); }
double stencil_comm_allreduce_max(double v) { return stencil::world().allreduce_max(v); }
void stencil_comm_finalize(void) { stencil::finalize_world(); }

int stencil_set_partition_ties(int xFirst) {
  const int prev = partition_tie_break() == stencil::detail::TieBreak::X ? 1 : 0;
  set_partition_tie_break(xFirst ? stencil::detail::TieBreak::X : stencil::detail::TieBreak::Z);
  return prev;
}

int stencil_rank_partition(const int64_t size[3], int64_t n, int64_t dim_out[3]) {
  if (n < 1) return -1;
  put(dim_out, RankPartition(d3(size), n).dim());
  return 0;
}
int stencil_rank_partition_subdomain(const int64_t size[3], int64_t n, const int64_t idx[3], int64_t size_out[3], int64_t origin_out[3]) {
  if (n < 1) return -1;
  RankPartition p(d3(size), n);
  if (!d3(idx).all_ge(0) || !d3(idx).all_lt(p.dim())) return -2;
  put(size_out, p.subdomain_size(d3(idx)));
  put(origin_out, p.subdomain_origin(d3(idx)));
  return 0;
}
int stencil_node_partition(const int64_t size[3], const uint64_t radius[27], int64_t nodes, int64_t gpus, int64_t sys_dim_out[3],
                           int64_t node_dim_out[3]) {
  if (nodes < 1 || gpus < 1) return -1;
  NodePartition p(d3(size), radius_from(radius), nodes, gpus);
  put(sys_dim_out, p.sys_dim());
  put(node_dim_out, p.node_dim());
  return 0;
}
int stencil_node_partition_subdomain(const int64_t size[3], const uint64_t radius[27], int64_t nodes, int64_t gpus, const int64_t idx[3],
                                     int64_t size_out[3], int64_t origin_out[3]) {
  if (nodes < 1 || gpus < 1) return -1;
  NodePartition p(d3(size), radius_from(radius), nodes, gpus);
  if (!d3(idx).all_ge(0) || !d3(idx).all_lt(p.dim())) return -2;
  put(size_out, p.subdomain_size(d3(idx)));
  put(origin_out, p.subdomain_origin(d3(idx)));
  return 0;
}
int stencil_qap_solve(const double *w, const double *d, int n, int exact, uint64_t *f_out, double *cost_out) {
  if (n < 1 || !w || !d || !f_out) return -1;
  Mat2D<double> W(n, n), D(n, n);
  std::memcpy(W.data(), w, sizeof(double) * size_t(n) * size_t(n));
  std::memcpy(D.data(), d, sizeof(double) * size_t(n) * size_t(n));
  double cost = 0;
  const std::vector<size_t> f = exact ? qap::solve(W, D, &cost) : qap::solve_catch(W, D, &cost);
  for (int i = 0; i < n; ++i) f_out[i] = f[size_t(i)];
  if (cost_out) *cost_out = cost;
  return 0;
}
int stencil_place_node_aware(const int64_t size[3], const uint64_t radius[27], int ngpus, const double *bandwidth, int *cuda_out,
                             int64_t dim_out[3]) {
  if (ngpus < 1 || !bandwidth || !cuda_out) return -1;
  std::vector<int> cudaOf(static_cast<size_t>(ngpus));
  for (int i = 0; i < ngpus; ++i) cudaOf[size_t(i)] = i;
  auto bw = [bandwidth, ngpus](int a, int b) { return bandwidth[a * ngpus + b]; };
  NodeAware p(d3(size), radius_from(radius), {{0}}, ngpus, cudaOf, bw);
  put(dim_out, p.dim());
  for (int i = 0; i < ngpus; ++i) cuda_out[i] = p.get_cuda(p.partition().idx(i));
  return 0;
}

double stencil_gpu_bandwidth(int src, int dst) { return gpu_topo::bandwidth(src, dst); }
int stencil_gpu_enable_peer(int src, int dst) {
  gpu_topo::enable_peer(src, dst);
  return gpu_topo::peer(src, dst) ? 1 : 0;
}

int stencil_pack(void *dst, const void *src, const int64_t src_pitch[3], const int64_t src_pos[3], const int64_t ext[3], size_t elem_size,
                 void *stream) {
  if (!dst || !src || !elem_size) return -1;
  stencil::pack(dst, src, d3(src_pitch), d3(src_pos), d3(ext), elem_size, st(stream));
  return 0;
}
int stencil_unpack(void *dst, const int64_t dst_pitch[3], const int64_t dst_pos[3], const int64_t ext[3], const void *src, size_t elem_size,
                   void *stream) {
  if (!dst || !src || !elem_size) return -1;
  stencil::unpack(dst, d3(dst_pitch), d3(dst_pos), d3(ext), src, elem_size, st(stream));
  return 0;
}
int stencil_translate(void *dst, const int64_t dst_pos[3], const int64_t dst_pitch[3], const void *src, const int64_t src_pos[3],
                      const int64_t src_pitch[3], const int64_t ext[3], size_t elem_size, void *stream) {
  if (!dst || !src || !elem_size) return -1;
  stencil::translate(dst, d3(dst_pos), d3(dst_pitch), src, d3(src_pos), d3(src_pitch), d3(ext), elem_size, st(stream));
  return 0;
}

int stencil_reduce_sum_f32(const float *src, const int64_t pitch[3], const int64_t ext[3], void *stream, double *sum_out) {
  if (!src || !sum_out) return -1;
  if (ext[0] < 0 || ext[1] < 0 || ext[2] < 0) return -2;
  *sum_out = stencil::reduce_sum_f32_sync(src, d3(pitch), d3(ext), st(stream));
  return 0;
}

/* ---- DistributedDomain ---- */
stencil_domain_t *stencil_domain_create(uint64_t x, uint64_t y, uint64_t z) {
  stencil_domain_t *h = new stencil_domain;
  h->owned.reset(new DistributedDomain(x, y, z));
  h->dd = h->owned.get();
  return h;
}
void stencil_domain_destroy(stencil_domain_t *dd) { delete dd; }
int stencil_domain_set_radius(stencil_domain_t *dd, const uint64_t radius[27]) {
  if (!dd) return -1;
  dd->dd->set_radius(radius_from(radius));
  return 0;
}
int stencil_domain_set_radius_constant(stencil_domain_t *dd, uint64_t r) {
  if (!dd) return -1;
  dd->dd->set_radius(size_t(r));
  return 0;
}
int stencil_domain_add_data(stencil_domain_t *dd, size_t elem_size, const char *name) {
  if (!dd || !elem_size) return -1;
  return int(dd->dd->add_data(elem_size, name ? name : ""));
}
int stencil_domain_set_methods(stencil_domain_t *dd, int methods) {
  if (!dd) return -1;
  dd->dd->set_methods(static_cast<MethodFlags>(methods));
  return 0;
}
int stencil_domain_set_placement(stencil_domain_t *dd, int placement) {
  if (!dd) return -1;
  dd->dd->set_placement(placement == STENCIL_PLACEMENT_TRIVIAL ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware);
  return 0;
}
int stencil_domain_set_gpus(stencil_domain_t *dd, const int *cuda_ids, int n) {
  if (!dd || n < 1 || !cuda_ids) return -1;
  dd->dd->set_gpus(std::vector<int>(cuda_ids, cuda_ids + n));
  return 0;
}
int stencil_domain_set_plan_file(stencil_domain_t *dd, int on) {
  if (!dd) return -1;
  dd->dd->set_plan_file(on != 0);
  return 0;
}
int stencil_domain_realize(stencil_domain_t *dd) {
  if (!dd) return -1;
  dd->dd->realize();
  return 0;
}
int stencil_domain_exchange(stencil_domain_t *dd) {
  if (!dd) return -1;
  dd->dd->exchange();
  return 0;
}
int stencil_domain_exchange_async(stencil_domain_t *dd) {
  if (!dd) return -1;
  dd->dd->exchange_async();
  return 0;
}
int stencil_domain_exchange_wait(stencil_domain_t *dd) {
  if (!dd) return -1;
  dd->dd->exchange_wait();
  return 0;
}
int stencil_domain_exchange_device_only(const stencil_domain_t *dd) { return dd ? (dd->dd->exchange_device_only() ? 1 : 0) : -1; }
void *stencil_domain_exchange_done_event(const stencil_domain_t *dd, int subdomain) {
  if (!dd || subdomain < 0 || size_t(subdomain) >= dd->dd->domains().size() || !dd->dd->exchange_device_only()) return nullptr;
  return dd->dd->exchange_done_event(size_t(subdomain));
}
int stencil_domain_exchange_after(stencil_domain_t *dd, int subdomain, void *cuda_event) {
  if (!dd || subdomain < 0 || size_t(subdomain) >= dd->dd->domains().size()) return -1;
  dd->dd->exchange_after(size_t(subdomain), static_cast<cudaEvent_t>(cuda_event));
  return 0;
}
int stencil_domain_swap(stencil_domain_t *dd) {
  if (!dd) return -1;
  dd->dd->swap();
  return 0;
}
int stencil_domain_num_subdomains(const stencil_domain_t *dd) { return dd ? int(dd->dd->domains().size()) : -1; }
int stencil_domain_grid(const stencil_domain_t *dd, int64_t grid_out[3]) {
  if (!dd || !grid_out) return -1;
  put(grid_out, dd->dd->subdomain_grid());
  return 0;
}
double stencil_domain_time_exchange(stencil_domain_t *dd, int iters, double *device_seconds) {
  return dd && iters > 0 ? dd->dd->time_exchange(iters, device_seconds) : -1.0;
}
uint64_t stencil_domain_exchange_bytes(const stencil_domain_t *dd, int methods) {
  return dd ? dd->dd->exchange_bytes_for_method(static_cast<MethodFlags>(methods)) : 0;
}
int stencil_domain_write_paraview(stencil_domain_t *dd, const char *prefix, int zero_nans) {
  if (!dd || !prefix) return -1;
  dd->dd->write_paraview(prefix, zero_nans != 0);
  return 0;
}
int stencil_subdomain_info(const stencil_domain_t *dd, int di, int64_t origin_out[3], int64_t size_out[3], int64_t raw_size_out[3],
                           int *gpu_out) {
  LocalDomain *ld = sub(dd, di);
  if (!ld) return -1;
  if (origin_out) put(origin_out, ld->origin());
  if (size_out) put(size_out, ld->size());
  if (raw_size_out) put(raw_size_out, ld->raw_size());
  if (gpu_out) *gpu_out = ld->gpu();
  return 0;
}
int stencil_subdomain_quantity(const stencil_domain_t *dd, int di, int qi, int64_t pitch_out[3], void **curr_out, void **next_out,
                               size_t *elem_size_out) {
  LocalDomain *ld = sub(dd, di);
  if (!ld || qi < 0 || qi >= ld->num_data()) return -1;
  if (pitch_out) put(pitch_out, ld->pitch(size_t(qi)));
  if (curr_out) *curr_out = ld->curr_data(size_t(qi));
  if (next_out) *next_out = ld->next_data(size_t(qi));
  if (elem_size_out) *elem_size_out = ld->elem_size(size_t(qi));
  return 0;
}
int stencil_subdomain_to_host(const stencil_domain_t *dd, int di, int qi, const int64_t pos[3], const int64_t ext[3], void *host_out) {
  LocalDomain *ld = sub(dd, di);
  if (!ld || qi < 0 || qi >= ld->num_data() || !host_out) return -1;
  const std::vector<unsigned char> v = ld->region_to_host(d3(pos), d3(ext), size_t(qi));
  std::memcpy(host_out, v.data(), v.size());
  return 0;
}
int stencil_subdomain_from_host(stencil_domain_t *dd, int di, int qi, const int64_t pos[3], const int64_t ext[3], const void *host_in) {
  LocalDomain *ld = sub(dd, di);
  if (!ld || qi < 0 || qi >= ld->num_data() || !host_in) return -1;
  ld->region_from_host(host_in, d3(pos), d3(ext), size_t(qi));
  return 0;
}
int stencil_subdomain_interior(const stencil_domain_t *dd, int di, int64_t lo_out[3], int64_t hi_out[3]) {
  if (!sub(dd, di)) return -1;
  const Rect3 r = dd->dd->get_interior()[size_t(di)];
  put(lo_out, r.lo);
  put(hi_out, r.hi);
  return 0;
}
int stencil_subdomain_exterior(const stencil_domain_t *dd, int di, int64_t lo_out[6][3], int64_t hi_out[6][3]) {
  if (!sub(dd, di)) return -1;
  const std::vector<Rect3> ex = dd->dd->get_exterior()[size_t(di)];
  for (size_t i = 0; i < ex.size() && i < 6; ++i) {
    put(lo_out[i], ex[i].lo);
    put(hi_out[i], ex[i].hi);
  }
  return int(ex.size());
}

/* ---- Jacobi3D ---- */
stencil_jacobi_t *stencil_jacobi_create(uint64_t x, uint64_t y, uint64_t z, int methods, int placement, int mode, const int *gpus, int ngpus,
                                        int tile_z) {
  stencil::Jacobi3D::Options opt;
  opt.methods = static_cast<MethodFlags>(methods);
  opt.placement = placement == STENCIL_PLACEMENT_TRIVIAL ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware;
  opt.mode = mode == STENCIL_JACOBI_OVERLAP ? stencil::Jacobi3D::Mode::Overlap
                                            : (mode == STENCIL_JACOBI_NO_OVERLAP ? stencil::Jacobi3D::Mode::NoOverlap : stencil::Jacobi3D::Mode::Fused);
  if (gpus && ngpus > 0) opt.gpus.assign(gpus, gpus + ngpus);
  if (tile_z > 0) opt.tileZ = tile_z;
  stencil_jacobi_t *h = new stencil_jacobi;
  h->j.reset(new stencil::Jacobi3D(x, y, z, opt));
  h->view.dd = &h->j->domain();
  return h;
}
void stencil_jacobi_destroy(stencil_jacobi_t *j) { delete j; }
int stencil_jacobi_init(stencil_jacobi_t *j) {
  if (!j) return -1;
  j->j->init();
  return 0;
}
int stencil_jacobi_step(stencil_jacobi_t *j) {
  if (!j) return -1;
  j->j->step();
  return 0;
}
int stencil_jacobi_sync(stencil_jacobi_t *j) {
  if (!j) return -1;
  j->j->sync();
  return 0;
}
double stencil_jacobi_run(stencil_jacobi_t *j, int iters) { return j ? j->j->run(iters) : -1.0; }
double stencil_jacobi_checksum(stencil_jacobi_t *j) { return j ? j->j->checksum() : 0.0; }
uint64_t stencil_jacobi_fingerprint(stencil_jacobi_t *j) { return j ? j->j->fingerprint() : 0; }
int stencil_jacobi_init_random(stencil_jacobi_t *j, unsigned seed) {
  if (!j) return -1;
  j->j->init_random(seed);
  return 0;
}
float stencil_jacobi_random_value(unsigned seed, int64_t x, int64_t y, int64_t z, int64_t gx, int64_t gy, int64_t gz) {
  return stencil::jacobi_random_value(seed, Dim3(x, y, z), Dim3(gx, gy, gz));
}
int stencil_jacobi_probe(stencil_jacobi_t *j, int64_t x, int64_t y, int64_t z, float *out) {
  if (!j || !out) return -1;
  return j->j->probe(Dim3(x, y, z), out) ? 1 : 0;
}
uint64_t stencil_jacobi_launches(const stencil_jacobi_t *j) { return j ? j->j->launches() : 0; }
int stencil_jacobi_mode(const stencil_jacobi_t *j) {
  if (!j) return -1;
  switch (j->j->mode()) {
  case stencil::Jacobi3D::Mode::Fused: return STENCIL_JACOBI_FUSED;
  case stencil::Jacobi3D::Mode::Overlap: return STENCIL_JACOBI_OVERLAP;
  default: return STENCIL_JACOBI_NO_OVERLAP;
  }
}
stencil_domain_t *stencil_jacobi_domain(stencil_jacobi_t *j) { return j ? &j->view : nullptr; }
int stencil_jacobi_step_io(stencil_jacobi_t *j, const void *params_pinned, size_t nbytes, float *result_pinned) {
  if (!j || !params_pinned || !result_pinned || nbytes > stencil::Jacobi3D::kParamBytes) return -1;
  j->j->step_io(params_pinned, nbytes, result_pinned);
  return 0;
}
int stencil_jacobi_host_step(float *next, const float *curr, int64_t x, int64_t y, int64_t z) {
  if (!next || !curr || x < 1 || y < 1 || z < 1) return -1;
  std::vector<float> c(curr, curr + x * y * z), n;
  stencil::jacobi_host_step(n, c, Dim3(x, y, z));
  std::memcpy(next, n.data(), n.size() * sizeof(float));
  return 0;
}

/* ---- AstarothSim ---- */
static stencil::AstarothSim::Stencil astaroth_kind(int k) {
  return k == STENCIL_ASTAROTH_STAR19_CROSS ? stencil::AstarothSim::Stencil::Star19Cross
                                            : (k == STENCIL_ASTAROTH_STAR19 ? stencil::AstarothSim::Stencil::Star19 : stencil::AstarothSim::Stencil::Average7);
}

stencil_astaroth_t *stencil_astaroth_create(uint64_t x, uint64_t y, uint64_t z, int methods, int placement, int stencil_kind, int fields,
                                            int fp64, int overlap, const int *gpus, int ngpus, double dt) {
  if (fields < 1) return nullptr;
  stencil::AstarothSim::Options opt;
  opt.methods = static_cast<MethodFlags>(methods);
  opt.placement = placement == STENCIL_PLACEMENT_TRIVIAL ? PlacementStrategy::Trivial : PlacementStrategy::NodeAware;
  opt.stencil = astaroth_kind(stencil_kind);
  opt.fields = fields;
  opt.fp64 = fp64 != 0;
  opt.autoSchedule = overlap < 0;
  opt.overlap = overlap != 0;
  if (dt > 0) opt.dt = dt;
  if (gpus && ngpus > 0) opt.gpus.assign(gpus, gpus + ngpus);
  stencil_astaroth_t *h = new stencil_astaroth;
  h->a.reset(new stencil::AstarothSim(x, y, z, opt));
  h->view.dd = &h->a->domain();
  return h;
}
void stencil_astaroth_destroy(stencil_astaroth_t *a) { delete a; }
int stencil_astaroth_init(stencil_astaroth_t *a) {
  if (!a) return -1;
  a->a->init();
  return 0;
}
int stencil_astaroth_step(stencil_astaroth_t *a) {
  if (!a) return -1;
  a->a->step();
  return 0;
}
int stencil_astaroth_sync(stencil_astaroth_t *a) {
  if (!a) return -1;
  a->a->sync();
  return 0;
}
double stencil_astaroth_run(stencil_astaroth_t *a, int iters, double *exchange_seconds) { return a ? a->a->run(iters, exchange_seconds) : -1.0; }
double stencil_astaroth_run_timed(stencil_astaroth_t *a, int iters, double *exchange_seconds, double *device_seconds) {
  return a ? a->a->run(iters, exchange_seconds, device_seconds) : -1.0;
}
uint64_t stencil_astaroth_launches(const stencil_astaroth_t *a) { return a ? a->a->launches() : 0; }
int stencil_astaroth_gather(stencil_astaroth_t *a, int field, double *out, uint64_t count) {
  if (!a || !out || field < 0 || field >= a->a->options().fields) return -1;
  const std::vector<double> v = a->a->gather(field);
  if (v.size() != count) return -2;
  std::memcpy(out, v.data(), v.size() * sizeof(double));
  return 0;
}
int stencil_astaroth_host_step(double *next, const double *curr, int64_t x, int64_t y, int64_t z, int stencil_kind, int fp64, double dt) {
  if (!next || !curr || x < 1 || y < 1 || z < 1) return -1;
  std::vector<double> c(curr, curr + x * y * z), n;
  stencil::astaroth_host_step(n, c, Dim3(x, y, z),
                              astaroth_kind(stencil_kind),
                              fp64 != 0, dt);
  std::memcpy(next, n.data(), n.size() * sizeof(double));
  return 0;
}
stencil_domain_t *stencil_astaroth_domain(stencil_astaroth_t *a) { return a ? &a->view : nullptr; }

} // extern "C"
