/* stencil_c.h — C ABI of libstencil_b200.so.

   The reference is a C++-only library; this flat ABI exists so that other runtimes (the Python package in
   stencil2_b200/, test drivers, Fortran/Julia codes) can use the same native objects through dlopen/ctypes.
   Conventions: every function returns 0 on success and a negative value on a usage error unless it returns a
   handle, a count or a value as documented; arrays of three are (x, y, z); radius arrays have 27 entries indexed
   ((z+1)*3 + (y+1))*3 + (x+1) for direction (x, y, z) in {-1,0,1}^3; `stream` arguments are cudaStream_t passed as
   void* (NULL = the legacy default stream). Fatal CUDA errors print a message and exit, as in the C++ API. */
#ifndef STENCIL_C_H
#define STENCIL_C_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---- */
/* "stencil-b200 <git rev> sm_100a" */
const char *stencil_version(void);
/* number of CUDA devices visible to this process (0 when there is no driver) */
int stencil_device_count(void);

/* page-locked host memory (cudaHostAlloc, portable): what per-step H2D/D2H copies should use */
void *stencil_host_alloc_pinned(size_t bytes);
void stencil_host_free_pinned(void *p);

/* ---- process group (reference: MPI_Init / MPI_Comm_rank / MPI_Finalize) ---- */
/* Join a TCP process group: rank 0 listens on addr:port. Replaces the environment-derived default. */
int stencil_comm_init_tcp(int rank, int size, const char *addr, int port);
int stencil_comm_rank(void);
int stencil_comm_size(void);
void stencil_comm_barrier(void);
double stencil_comm_allreduce_max(double v);
void stencil_comm_finalize(void);

/* ---- partition, QAP and placement: pure CPU (reference: partition.hpp, qap.hpp) ---- */
/* which axis is cut when the criterion ties: 0 = z, then y, then x (default; keeps x rows long and faces contiguous),
   1 = x first (the reference's order). Returns the previous setting. Process-wide; set before realize(). */
int stencil_set_partition_ties(int xFirst);
int stencil_rank_partition(const int64_t size[3], int64_t n, int64_t dim_out[3]);
int stencil_rank_partition_subdomain(const int64_t size[3], int64_t n, const int64_t idx[3], int64_t size_out[3],
                                     int64_t origin_out[3]);
int stencil_node_partition(const int64_t size[3], const uint64_t radius[27], int64_t nodes, int64_t gpus,
                           int64_t sys_dim_out[3], int64_t node_dim_out[3]);
int stencil_node_partition_subdomain(const int64_t size[3], const uint64_t radius[27], int64_t nodes, int64_t gpus,
                                     const int64_t idx[3], int64_t size_out[3], int64_t origin_out[3]);
/* w and d are n*n row-major. exact != 0: exhaustive search, else 2-swap local search. f_out[n]; cost_out optional. */
int stencil_qap_solve(const double *w, const double *d, int n, int exact, uint64_t *f_out, double *cost_out);
/* Node-aware placement of n = ngpus subdomains of one node given a bandwidth matrix (n*n, row-major, CUDA ids 0..n-1).
   cuda_out[i] = device of subdomain i (linear order, x fastest); dim_out = subdomain grid. */
int stencil_place_node_aware(const int64_t size[3], const uint64_t radius[27], int ngpus, const double *bandwidth,
                             int *cuda_out, int64_t dim_out[3]);

/* ---- GPU topology (reference: gpu_topology.hpp) ---- */
double stencil_gpu_bandwidth(int src, int dst);
int stencil_gpu_enable_peer(int src, int dst); /* returns 1 if peer access is (now) enabled */

/* ---- standalone box copies on the current device (reference: pack_kernel.cuh, copy.cuh) ---- */
int stencil_pack(void *dst, const void *src, const int64_t src_pitch[3], const int64_t src_pos[3], const int64_t ext[3],
                 size_t elem_size, void *stream);
int stencil_unpack(void *dst, const int64_t dst_pitch[3], const int64_t dst_pos[3], const int64_t ext[3], const void *src,
                   size_t elem_size, void *stream);
int stencil_translate(void *dst, const int64_t dst_pos[3], const int64_t dst_pitch[3], const void *src,
                      const int64_t src_pos[3], const int64_t src_pitch[3], const int64_t ext[3], size_t elem_size,
                      void *stream);

/* ---- reductions (no counterpart in the reference) ---- */
/* Sum (fp64 accumulation, order-independent result) of a box of an fp32 pitched device array on the current
   device. Allocates its scratch and result on the device, synchronises `stream`, frees, and writes *sum_out. */
int stencil_reduce_sum_f32(const float *src, const int64_t pitch[3], const int64_t ext[3], void *stream, double *sum_out);

/* Allocation-free form of stencil_reduce_sum_f32 (no counterpart in the reference).
   For callers that must not call cudaMalloc/cudaFree where they need the sum: a monitor inside a free-running
   fused iteration loop (allocation and free synchronise the device and would stall GPUs that wait on each other's
   flags), a solver whose subdomains fill HBM, or code living under somebody else's memory pool. The reduction's
   scratch (per-CTA partials + a ticket) is a fixed device array inside the library, STENCIL_NOALLOC_SCRATCH_BYTES
   per GPU, held from library load to process exit. The function calls no allocator (CUDA's, torch's or any other)
   and does not synchronise:
     src, pitch, ext   as stencil_reduce_sum_f32, on the CURRENT device
     sum_out           where the kernel writes the fp64 sum: device memory or mapped pinned host memory, caller-owned
     stream            the result is valid once work enqueued on `stream` up to this call has completed
   Limits: ext[0] <= STENCIL_NOALLOC_MAX_ROW_CELLS and ext[1]*ext[2] <= STENCIL_NOALLOC_MAX_ROWS; larger boxes are
   refused with a message on stderr and return value -3 (nothing is launched). Other errors: -1 null argument,
   -2 inconsistent pitch/extent, -10..-12 CUDA set-up failures.
   The FIRST call on each device looks up the scratch array's device address and caches it — make that first call
   outside any stream capture. All calls on a device share the one scratch array, which keeps the previous call's
   contents: the ticket is cleared (stream-ordered) at every call, but two calls whose kernels can overlap on the
   GPU — different streams without an ordering between them — race on it and give wrong sums. Serialise them. */
#define STENCIL_NOALLOC_SCRATCH_BYTES 9728
#define STENCIL_NOALLOC_MAX_ROW_CELLS 2147483647LL
#define STENCIL_NOALLOC_MAX_ROWS 2147483647LL
int stencil_reduce_sum_f32_noalloc(const float *src, const int64_t pitch[3], const int64_t ext[3], void *stream, double *sum_out);

/* ---- DistributedDomain (reference: stencil.hpp) ---- */
typedef struct stencil_domain stencil_domain_t;
enum {
  STENCIL_METHOD_NONE = 0,
  STENCIL_METHOD_CUDA_MPI = 1,
  STENCIL_METHOD_CUDA_AWARE_MPI = 2,
  STENCIL_METHOD_COLOCATED = 4,
  STENCIL_METHOD_MEMCPY_PEER = 8,
  STENCIL_METHOD_KERNEL = 16,
  STENCIL_METHOD_NVLINK_DIRECT = 32,
  STENCIL_METHOD_ALL = 1 + 4 + 8 + 16 + 32
};
enum { STENCIL_PLACEMENT_NODE_AWARE = 0, STENCIL_PLACEMENT_TRIVIAL = 1 };

stencil_domain_t *stencil_domain_create(uint64_t x, uint64_t y, uint64_t z);
void stencil_domain_destroy(stencil_domain_t *dd);
int stencil_domain_set_radius(stencil_domain_t *dd, const uint64_t radius[27]);
int stencil_domain_set_radius_constant(stencil_domain_t *dd, uint64_t r);
/* returns the quantity index */
int stencil_domain_add_data(stencil_domain_t *dd, size_t elem_size, const char *name);
int stencil_domain_set_methods(stencil_domain_t *dd, int methods);
int stencil_domain_set_placement(stencil_domain_t *dd, int placement);
int stencil_domain_set_gpus(stencil_domain_t *dd, const int *cuda_ids, int n);
int stencil_domain_set_plan_file(stencil_domain_t *dd, int on);
int stencil_domain_realize(stencil_domain_t *dd);
int stencil_domain_exchange(stencil_domain_t *dd);
/* the two halves of exchange(): enqueue the device work / drive host-mediated transports and wait (see
   DistributedDomain::exchange_async in stencil/stencil.hpp). When every message travels by device stores or copies
   (stencil_domain_exchange_device_only() == 1) the host never has to wait: make your stream wait on
   stencil_domain_exchange_done_event(dd, i) (a cudaEvent_t, NULL otherwise) and hand the event that ends your use of
   subdomain i to stencil_domain_exchange_after() before the next stencil_domain_exchange_async(). */
int stencil_domain_exchange_async(stencil_domain_t *dd);
int stencil_domain_exchange_wait(stencil_domain_t *dd);
int stencil_domain_exchange_device_only(const stencil_domain_t *dd);
void *stencil_domain_exchange_done_event(const stencil_domain_t *dd, int subdomain);
int stencil_domain_exchange_after(stencil_domain_t *dd, int subdomain, void *cuda_event);
int stencil_domain_swap(stencil_domain_t *dd);
int stencil_domain_num_subdomains(const stencil_domain_t *dd);
/* how the domain was cut: subdomains along x, y, z over the whole job (valid after realize) */
int stencil_domain_grid(const stencil_domain_t *dd, int64_t grid_out[3]);
/* `iters` blocking exchanges between barriers: returns host wall seconds (max over ranks); *device_seconds (may be
   NULL) = time between CUDA events on every exchange stream around the calls (max over GPUs and ranks).
   Reference: the loop bin/weak_exchange.cu times. */
double stencil_domain_time_exchange(stencil_domain_t *dd, int iters, double *device_seconds);
uint64_t stencil_domain_exchange_bytes(const stencil_domain_t *dd, int methods);
int stencil_domain_write_paraview(stencil_domain_t *dd, const char *prefix, int zero_nans);
/* geometry of local subdomain di */
int stencil_subdomain_info(const stencil_domain_t *dd, int di, int64_t origin_out[3], int64_t size_out[3],
                           int64_t raw_size_out[3], int *gpu_out);
/* pitch (elements) and device pointers to logical element (0,0,0) of quantity qi */
int stencil_subdomain_quantity(const stencil_domain_t *dd, int di, int qi, int64_t pitch_out[3], void **curr_out,
                               void **next_out, size_t *elem_size_out);
/* dense host copies of a box given as offset/extent (elements) inside the allocation incl. halo */
int stencil_subdomain_to_host(const stencil_domain_t *dd, int di, int qi, const int64_t pos[3], const int64_t ext[3], void *host_out);
int stencil_subdomain_from_host(stencil_domain_t *dd, int di, int qi, const int64_t pos[3], const int64_t ext[3], const void *host_in);
/* interior box and up to 6 exterior slabs, each as lo[3] hi[3] (global coordinates); returns the number of slabs */
int stencil_subdomain_interior(const stencil_domain_t *dd, int di, int64_t lo_out[3], int64_t hi_out[3]);
int stencil_subdomain_exterior(const stencil_domain_t *dd, int di, int64_t lo_out[6][3], int64_t hi_out[6][3]);

/* ---- Jacobi3D engine (reference: bin/jacobi3d.cu) ---- */
typedef struct stencil_jacobi stencil_jacobi_t;
enum { STENCIL_JACOBI_FUSED = 0, STENCIL_JACOBI_OVERLAP = 1, STENCIL_JACOBI_NO_OVERLAP = 2 };
/* gpus may be NULL (ngpus = 0) for the rank's default devices */
stencil_jacobi_t *stencil_jacobi_create(uint64_t x, uint64_t y, uint64_t z, int methods, int placement, int mode,
                                        const int *gpus, int ngpus, int tile_z);
void stencil_jacobi_destroy(stencil_jacobi_t *j);
int stencil_jacobi_init(stencil_jacobi_t *j);
/* initial field = pseudo-random values in [0,1) that depend only on seed and the global coordinate of the cell */
int stencil_jacobi_init_random(stencil_jacobi_t *j, unsigned seed);
float stencil_jacobi_random_value(unsigned seed, int64_t x, int64_t y, int64_t z, int64_t gx, int64_t gy, int64_t gz);
int stencil_jacobi_step(stencil_jacobi_t *j);
int stencil_jacobi_sync(stencil_jacobi_t *j);
/* device-timed seconds for `iters` iterations: CUDA events per local GPU, max over GPUs and ranks */
double stencil_jacobi_run(stencil_jacobi_t *j, int iters);
double stencil_jacobi_checksum(stencil_jacobi_t *j);
/* 64-bit fingerprint of the whole field (all ranks): independent of partition and schedule, equal exactly when every
   cell agrees bit for bit (up to a 2^-64 accident) */
uint64_t stencil_jacobi_fingerprint(stencil_jacobi_t *j);
/* returns 1 and writes *out if this rank owns the cell, else 0 */
int stencil_jacobi_probe(stencil_jacobi_t *j, int64_t x, int64_t y, int64_t z, float *out);
uint64_t stencil_jacobi_launches(const stencil_jacobi_t *j);
int stencil_jacobi_mode(const stencil_jacobi_t *j);
stencil_domain_t *stencil_jacobi_domain(stencil_jacobi_t *j); /* borrowed: do not destroy */
/* One iteration with per-step host I/O, stream-ordered with the step (no host blocking): copies `nbytes` (<= 64)
   of run-time parameters from PINNED host memory to every local GPU (floats: [0] hot, [1] cold sphere temperature),
   runs the step, then copies the new value at the centre of local subdomain di to result_pinned[di] (pinned).
   Call stencil_jacobi_sync() before reading the results. */
int stencil_jacobi_step_io(stencil_jacobi_t *j, const void *params_pinned, size_t nbytes, float *result_pinned);
/* host reference: one iteration on a dense periodic x*y*z fp32 array (same arithmetic as the kernel) */
int stencil_jacobi_host_step(float *next, const float *curr, int64_t x, int64_t y, int64_t z);

/* ---- AstarothSim engine (reference: bin/astaroth_sim.cu): radius 3, `fields` quantities of fp32 or fp64 ---- */
typedef struct stencil_astaroth stencil_astaroth_t;
#define STENCIL_ASTAROTH_AVERAGE7 0 /* the reference's stand-in compute */
#define STENCIL_ASTAROTH_STAR19 1   /* next = curr + dt * (6th-order 19-point Laplacian) */
#define STENCIL_ASTAROTH_STAR19_CROSS 2 /* ... + 0.25 * (6th-order mixed second derivatives + a corner term): reads all 26 halo directions */
/* overlap: 1 interior/exchange/exterior schedule, 0 exchange then one whole-region launch, -1 engine's choice */
stencil_astaroth_t *stencil_astaroth_create(uint64_t x, uint64_t y, uint64_t z, int methods, int placement, int stencil_kind, int fields,
                                            int fp64, int overlap, const int *gpus, int ngpus, double dt);
void stencil_astaroth_destroy(stencil_astaroth_t *a);
int stencil_astaroth_init(stencil_astaroth_t *a);
int stencil_astaroth_step(stencil_astaroth_t *a);
int stencil_astaroth_sync(stencil_astaroth_t *a);
/* `iters` iterations; host wall seconds (max over ranks); *exchange_seconds (may be NULL) = time inside exchange() */
double stencil_astaroth_run(stencil_astaroth_t *a, int iters, double *exchange_seconds);
/* same, and *device_seconds = time between CUDA events on the compute streams around the loop (max over GPUs and ranks) */
double stencil_astaroth_run_timed(stencil_astaroth_t *a, int iters, double *exchange_seconds, double *device_seconds);
uint64_t stencil_astaroth_launches(const stencil_astaroth_t *a);
/* dense [z][y][x] copy of one field as double into out[x*y*z]; cells owned by other ranks are NaN */
int stencil_astaroth_gather(stencil_astaroth_t *a, int field, double *out, uint64_t count);
/* host reference of one step on a dense periodic array */
int stencil_astaroth_host_step(double *next, const double *curr, int64_t x, int64_t y, int64_t z, int stencil_kind, int fp64, double dt);
stencil_domain_t *stencil_astaroth_domain(stencil_astaroth_t *a); /* borrowed: do not destroy */

#ifdef __cplusplus
}
#endif
#endif /* STENCIL_C_H */
