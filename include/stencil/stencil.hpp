#pragma once
/* DistributedDomain: the user-facing object. Configure -> realize() -> { compute; exchange(); swap(); } loop.

   Public surface follows the reference (include/stencil/stencil.hpp:29-354, src/stencil.cu:6-939): MethodFlags,
   PlacementStrategy, set_radius/add_data/set_methods/set_placement/set_gpus, realize, exchange, swap,
   get_interior/get_exterior, get_compute_region, get_origin, exchange_bytes_for_method, write_paraview, the public
   timing fields, plan_<rank>.txt.

   What is different underneath (B200 / NVSwitch first):
   * a new method, MethodFlags::NvlinkDirect (part of All): every message whose receiver sits on a peer-accessible
     GPU of this node — same process or another — is a segment of ONE device-initiated push per subdomain
     (src/box_copy.cu): boundary cells are stored straight into the neighbour's halo cells through NVLink, and the
     last CTA publishes an epoch flag in the neighbour's slab (fence.sys + store). No pack buffer, no cudaMemcpyPeer,
     no unpack kernel, no host in the data path. Receivers wait with one acquire-spin warp (other GPU) or a stream
     event (same GPU). The legacy flags keep their meaning and select legacy-equivalent paths:
       CudaKernel        same rank + same GPU        -> local segments of the push plan
       CudaMemcpyPeer    same rank, peer GPU         -> pack, cudaMemcpyPeerAsync, event, unpack
       CudaMpiColocated  other rank on node, peer    -> pack, cudaMemcpyPeerAsync into an IPC-mapped buffer, IPC event
       CudaMpi           anything (incl. self)       -> pack, D2H pinned, Comm send/recv, H2D, unpack  ("staged")
       CudaAwareMpi      only with STENCIL_USE_CUDA_AWARE_MPI (needs a real MPI under the Comm)
     priority: Kernel > NvlinkDirect > MemcpyPeer > Colocated > CudaAwareMpi > CudaMpi.
   * process group = stencil::Comm (comm.hpp), not MPI: one process may drive all GPUs (reference default) or
     `torchrun`-style one process per GPU; peers of other processes are mapped with CUDA IPC.
   * swap() is O(1) host work. */
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "stencil/boundary.hpp"
#include "stencil/comm.hpp"
#include "stencil/dim3.hpp"
#include "stencil/gpu_topology.hpp"
#include "stencil/local_domain.cuh"
#include "stencil/logging.hpp"
#include "stencil/mpi_topology.hpp"
#include "stencil/partition.hpp"
#include "stencil/radius.hpp"
#include "stencil/rcstream.hpp"
#include "stencil/rect3.hpp"
#include "stencil/tx_common.hpp"

#ifndef STENCIL_USE_CUDA_AWARE_MPI
#define STENCIL_USE_CUDA_AWARE_MPI 0
#endif

enum class MethodFlags : int {
  None = 0,
  CudaMpi = 1,
  CudaAwareMpi = 2,
  CudaMpiColocated = 4,
  CudaMemcpyPeer = 8,
  CudaKernel = 16,
  NvlinkDirect = 32,
#if STENCIL_USE_CUDA_AWARE_MPI == 1
  All = 1 + 2 + 4 + 8 + 16 + 32
#else
  All = 1 + 4 + 8 + 16 + 32
#endif
};
static_assert(sizeof(MethodFlags) == sizeof(int), "int");

inline MethodFlags operator|(MethodFlags a, MethodFlags b) { return static_cast<MethodFlags>(static_cast<int>(a) | static_cast<int>(b)); }
inline MethodFlags &operator|=(MethodFlags &a, MethodFlags b) { return a = a | b; }
inline MethodFlags operator&(MethodFlags a, MethodFlags b) { return static_cast<MethodFlags>(static_cast<int>(a) & static_cast<int>(b)); }
/* true if the two sets share a method (reference spelling) */
inline bool operator&&(MethodFlags a, MethodFlags b) { return (a & b) != MethodFlags::None; }
inline bool any(MethodFlags a) { return a != MethodFlags::None; }
/* "staged/colo/peer/kernel/direct" style label used in CSV output */
std::string method_string(MethodFlags m, const char *sep = "/");

namespace stencil {
class Exchanger; // owns plans, transports, streams (src/stencil.cu)

/* What a device-initiated (fused compute + push) kernel needs to know about the subdomain on side `dir` of a local
   subdomain: a strided view of one of its quantities as seen from the local GPU (a peer / IPC-mapped address when
   it lives elsewhere), its control-block flag words, and where it runs. `reachable` is false when the neighbour
   is on another node or peer access is unavailable (then only exchange() can talk to it). */
struct NeighborInfo {
  bool exists = false;    // false across a non-periodic global boundary
  bool reachable = false; // loads/stores from the local GPU work
  bool sameRank = false, sameDevice = false;
  int rank = -1, dom = -1, dev = -1;
  Dim3 idx, size;
  BoxView view;             // base[b] = buffer b (0/1) of the quantity, logical element (0,0,0)
  unsigned *control = nullptr; // the neighbour's control block (LocalDomain::kControlBytes), as mapped locally
  char *slab = nullptr;               // base of the neighbour's slab as mapped locally
  const LocalDomain *layout = nullptr; // its geometry/layout (a shadow object for other ranks): slab offsets, mailboxes
};
} // namespace stencil

class DistributedDomain {
  friend class stencil::Exchanger;

public:
  DistributedDomain(size_t x, size_t y, size_t z);
  /* use an explicit process group instead of stencil::world() */
  DistributedDomain(size_t x, size_t y, size_t z, stencil::Comm &comm);
  ~DistributedDomain();
  DistributedDomain(const DistributedDomain &) = delete;
  DistributedDomain &operator=(const DistributedDomain &) = delete;

  const Dim3 &size() const noexcept { return size_; }
  std::vector<LocalDomain> &domains() { return domains_; }
  const std::vector<LocalDomain> &domains() const { return domains_; }

  void set_radius(size_t r) { radius_ = Radius::constant(r); }
  void set_radius(const Radius &r) { radius_ = r; }
  const Radius &radius() const noexcept { return radius_; }

  template <typename T> DataHandle<T> add_data(const std::string &name = "") {
    elemSizes_.push_back(sizeof(T));
    names_.push_back(name);
    return DataHandle<T>(elemSizes_.size() - 1, name);
  }
  /* untyped quantity of n bytes per element */
  size_t add_data(size_t n, const std::string &name) {
    elemSizes_.push_back(n);
    names_.push_back(name);
    return elemSizes_.size() - 1;
  }

  void set_methods(MethodFlags f) noexcept { flags_ = f; }
  bool any_methods(MethodFlags f) const noexcept { return (flags_ & f) != MethodFlags::None; }
  MethodFlags methods() const noexcept { return flags_; }
  void set_placement(PlacementStrategy s) noexcept { strategy_ = s; }

  /* Choose the CUDA devices this rank contributes (one subdomain per entry; repeat an id to oversubscribe).
     README of the reference calls this use_gpus; both spellings work. */
  void set_gpus(const std::vector<int> &cudaIds);
  void use_gpus(const std::vector<int> &cudaIds) { set_gpus(cudaIds); }
  const std::vector<int> &gpus() const noexcept { return gpus_; }

  /* Optional: non-periodic global boundaries (the reference is periodic-only). */
  void set_boundary(const Boundary &b) { boundary_ = b; }
  /* Optional: row alignment of the local arrays in bytes (default 128). */
  void set_alignment(size_t bytes) { align_ = bytes; }
  /* Optional: bandwidth(cudaA, cudaB) model for NodeAware placement (default: gpu_topo::bandwidth). */
  void set_bandwidth_model(NodeAware::BandwidthFn f) { bandwidthFn_ = std::move(f); }

  const Dim3 &get_origin(int64_t i) const { return domains_.at(size_t(i)).origin(); }
  const Rect3 get_compute_region() const noexcept { return Rect3(Dim3(0, 0, 0), size_); }
  /* index of local subdomain i in the global subdomain grid, and the grid itself */
  Dim3 get_subdomain_idx(int64_t i) const;
  Dim3 subdomain_grid() const;

  /* bytes sent per exchange() over the whole job by the given method(s) */
  uint64_t exchange_bytes_for_method(const MethodFlags &m) const;

  void realize();
  /* O(1): flips curr/next of every local subdomain */
  void swap();
  /* Blocking halo exchange of all `curr` buffers: returns when every local halo is filled and every local send has
     been consumed. Launch interior work on your own streams before calling to overlap. */
  void exchange();

  /* The same exchange in two halves, for callers that overlap it with their own work or never block the host:
       exchange_async()  enqueues all device work of the exchange on the exchange streams and returns;
       exchange_wait()   drives the host-mediated transports (staged / colocated) and waits for the streams.
     exchange() == exchange_async(); exchange_wait();
     If exchange_device_only() (every message travels by device stores or copies — always the case on one NVSwitch box
     with the default methods), the host never has to wait: make your stream wait on exchange_done_event(i) before it
     reads the halos of subdomain i, and hand the event that ends your use of subdomain i's buffers (the kernels that
     wrote what will be sent next AND the kernels that read the halos about to be overwritten) to
     exchange_after(i, e) before the next exchange_async(). Flag time-outs are reported by the next call. */
  void exchange_async();
  void exchange_wait();
  bool exchange_device_only() const;
  cudaEvent_t exchange_done_event(size_t i) const;
  void exchange_after(size_t i, cudaEvent_t e);
  /* For engines whose compute kernel writes the +-x face halo cells of same-GPU neighbours itself (the strided
     faces are the expensive part of an exchange, and the kernel has those values in registers): leave those messages
     out of exchange() from now on / put them back. Everything else — other GPUs, edges, corners, y and z faces — is
     still exchanged. exchange_x_faces_all_local(): true if that covers every +-x face message of this rank. */
  void exchange_skip_local_x_faces(bool on);
  bool exchange_x_faces_all_local() const;

  /* regions whose stencil does not read halo cells / up to six slabs that do (order +x,+y,+z,-x,-y,-z) */
  std::vector<Rect3> get_interior() const;
  std::vector<std::vector<Rect3>> get_exterior() const;

  /* CSV dump "<prefix>_<id>.txt": header Z,Y,X,<names>; one row per compute cell (4- and 8-byte quantities) */
  void write_paraview(const std::string &prefix, bool zeroNaNs = false);

  int rank() const noexcept { return rank_; }
  int world_size() const noexcept { return worldSize_; }
  stencil::Comm &comm() const noexcept { return *comm_; }
  Placement *placement() const noexcept { return placement_.get(); }
  /* the stream exchange() uses for local subdomain i (HIGH priority, on that subdomain's GPU) */
  cudaStream_t exchange_stream(size_t i) const;
  /* `iters` blocking exchange() calls between barriers: returns the host wall seconds (max over ranks) and, in
     *deviceSeconds (may be null), the device time between CUDA events recorded on every exchange stream before the
     first and after the last call (max over local GPUs and ranks). What bin/weak_exchange.cu prints. */
  double time_exchange(int iters, double *deviceSeconds = nullptr);
  /* number of completed exchange() calls (the epoch published in the arrival flags) */
  unsigned epoch() const noexcept;
  stencil::Exchanger *exchanger() const noexcept { return ex_.get(); }
  /* neighbour of local subdomain di on side dir, for quantity qi (valid after realize()) */
  stencil::NeighborInfo neighbor(size_t di, const Dim3 &dir, size_t qi);

  /* accumulated max-over-ranks seconds; valid on rank 0 (exchange/swap only with STENCIL_EXCHANGE_STATS or
     set_exchange_stats(true): they add a barrier per call) */
  double timeExchange_ = 0, timeSwap_ = 0;
  double timeMpiTopo_ = 0, timeNodeGpus_ = 0, timePeerEn_ = 0, timePlacement_ = 0, timePlan_ = 0, timeRealize_ = 0,
         timeCreate_ = 0;
  void set_exchange_stats(bool on) noexcept { exchangeStats_ = on; }
  /* write plan_<rank>.txt at realize() (default on, as the reference) */
  void set_plan_file(bool on) noexcept { planFile_ = on; }

private:
  void init(size_t x, size_t y, size_t z);

  Dim3 size_;
  stencil::Comm *comm_;
  int rank_ = 0, worldSize_ = 1;
  std::unique_ptr<MpiTopology> topo_;
  std::vector<int> gpus_;
  Radius radius_;
  Boundary boundary_;
  size_t align_ = 128;
  MethodFlags flags_ = MethodFlags::All;
  PlacementStrategy strategy_ = PlacementStrategy::NodeAware;
  NodeAware::BandwidthFn bandwidthFn_;
  std::vector<size_t> elemSizes_;
  std::vector<std::string> names_;
  std::unique_ptr<Placement> placement_;
  std::vector<LocalDomain> domains_;
  std::unique_ptr<stencil::Exchanger> ex_;
  bool exchangeStats_ =
#ifdef STENCIL_EXCHANGE_STATS
      true;
#else
      false;
#endif
  bool planFile_ = true;
  uint64_t numBytesCudaMpi_ = 0, numBytesCudaMpiColocated_ = 0, numBytesCudaMemcpyPeer_ = 0, numBytesCudaKernel_ = 0,
           numBytesNvlinkDirect_ = 0, numBytesCudaAwareMpi_ = 0;
};
