/* DistributedDomain + Exchanger: placement, allocation, exchange planning and the per-iteration exchange.
   Reference: include/stencil/stencil.hpp:133-255 (ctor), src/stencil.cu:27-539 (realize), :541-666 (swap, interior /
   exterior), :670-864 (exchange), :866-939 (write_paraview). See stencil.hpp for the design. */
#include "stencil/stencil.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <tuple>

#include "stencil/copy_plan.hpp"
#include "stencil/detail/launch_pool.hpp"
#include "stencil/nvtx.hpp"
#include "stencil/packer.cuh"

std::string method_string(MethodFlags m, const char *sep) {
  std::string s;
  auto add = [&](MethodFlags f, const char *name) {
    if (m && f) {
      if (!s.empty()) s += sep;
      s += name;
    }
  };
  add(MethodFlags::CudaMpi, "staged");
  add(MethodFlags::CudaAwareMpi, "cuda-aware");
  add(MethodFlags::CudaMpiColocated, "colo");
  add(MethodFlags::CudaMemcpyPeer, "peer");
  add(MethodFlags::CudaKernel, "kernel");
  add(MethodFlags::NvlinkDirect, "direct");
  return s;
}

namespace stencil {

namespace {

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int d) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != d) CUDA_RUNTIME(cudaSetDevice(d));
  }
  ~DevGuard() { cudaSetDevice(prev); }
};

enum class Route { Kernel, Direct, PeerCopy, Colo, Staged };

const char *route_name(Route r) {
  switch (r) {
  case Route::Kernel: return "kernel";
  case Route::Direct: return "direct";
  case Route::PeerCopy: return "peer";
  case Route::Colo: return "colo";
  default: return "staged";
  }
}

/* where a subdomain lives */
struct Where {
  Dim3 idx;
  int rank, dom, dev;
};

/* tags of the exchange's own host messages; kinds 0-3 are the reference's make_tag kinds (IPC set-up + notify) */
int staged_tag(int srcDom, int dstDom) { return (1 << 24) | ((srcDom & 0xFF) << 8) | (dstDom & 0xFF); }
int ack_tag(int payload) { return (2 << 24) | (payload & 0xFFFF); }

} // namespace

unsigned long long flag_timeout_ns() {
  const char *e = std::getenv("STENCIL_FLAG_TIMEOUT_MS");
  const double ms = e ? std::atof(e) : 5000.0;
  return (unsigned long long)((ms > 0 ? ms : 5000.0) * 1e6);
}

/* ------------------------------------------------------------------------------------------------------------ */

class Exchanger {
  DistributedDomain &dd_;
  const size_t nd_; // local subdomains

  std::vector<RcStream> streams_;
  std::vector<cudaEvent_t> pushDone_;
  std::vector<cudaEvent_t> exDone_; // end of the enqueued exchange on each subdomain's stream (device-only plans)
  std::vector<cudaEvent_t> after_;  // caller's events the next enqueue() waits for (not owned)
  bool inFlight_ = false;
  /* STENCIL_EXCHANGE_TRACE=n: print per-subdomain device times of the first n exchanges (push kernel, flag wait) */
  int trace_ = 0;
  std::vector<cudaEvent_t> trA_, trB_, trC_;
  unsigned epoch_ = 0;
  /* several local GPUs: the per-subdomain parts of enqueue() run on one host thread per subdomain */
  stencil::detail::LaunchPool pool_;
  bool threads_ = true;
  void for_each_dom(const std::function<void(size_t)> &fn) {
    if (threads_ && nd_ > 1)
      pool_.run(nd_, fn);
    else
      for (size_t di = 0; di < nd_; ++di) fn(di);
  }
  unsigned *errWord_ = nullptr; // pinned + mapped
  unsigned long long timeoutNs_;

  /* ---- direct push (Kernel + NvlinkDirect) ---- */
  std::vector<CopyPlan> push_;
  /* the +-x FACE messages that stay on the sender's GPU (wrap onto itself, or an oversubscribed neighbour) are optional
     segments of the push plan: an engine whose compute kernel writes those halo cells itself (AstarothSim) turns them off */
  bool skipXLocal_ = false;
  size_t xFaceMsgs_ = 0, xFaceLocal_ = 0;
  std::vector<CopyPlan> scatter_; // after the flags: my x-side mailboxes -> my halo cells (thin rows from other GPUs)
  std::vector<SignalSet> arriveSig_; // flags I raise in cross-device neighbours after my push
  std::vector<WaitSet> arriveWait_;  // my flags they raise
  std::vector<SignalSet> readySig_;  // cross-PROCESS neighbours: "my halos may be overwritten for epoch e"
  std::vector<WaitSet> readyWait_;
  std::vector<std::vector<int>> sameDevDeps_; // local subdomains on my GPU whose push must finish before I return

  /* ---- IPC mapping of other ranks' slabs ---- */
  struct RemoteSlab {
    void *ptr;
    std::unique_ptr<LocalDomain> shadow; // layout only
  };
  std::map<std::tuple<int, int, int>, RemoteSlab> remote_; // (my device, rank, dom)
  std::vector<cudaIpcMemHandle_t> allHandles_;             // indexed by slabIndex(rank, dom)
  std::vector<int> domOffsetOfRank_;

  /* ---- peer copy (legacy CudaMemcpyPeer) ---- */
  struct PeerCopy {
    int src, dst;
    std::unique_ptr<DevicePacker> packer;
    std::unique_ptr<DeviceUnpacker> unpacker;
    cudaEvent_t copied;
  };
  std::vector<PeerCopy> peerCopies_;

  /* ---- staged (legacy CudaMpi) ---- */
  struct StagedSend {
    int src, dstRank, tag;
    std::unique_ptr<DevicePacker> packer;
    void *host = nullptr; // pinned staging buffer; unused when the communicator takes device pointers
    cudaEvent_t staged;
    bool pending = false;
  };
  /* CudaAwareMpi (reference C29, tx_cuda.cuh:820-990): the packed device buffer itself is handed to the
     communicator, no pinned host copy on either side. Needs Comm::device_buffers(). */
  bool deviceMsgs_ = false;
  struct StagedRecv {
    int dst, srcRank, tag;
    std::unique_ptr<DeviceUnpacker> unpacker;
    void *host = nullptr;
    bool pending = false;
  };
  std::vector<StagedSend> stagedSends_;
  std::vector<StagedRecv> stagedRecvs_;

  /* ---- colocated (legacy CudaMpiColocated): pack -> peer copy into the receiver's IPC-mapped buffer ---- */
  struct ColoSend {
    int src, dstRank, tagBase;
    std::unique_ptr<DevicePacker> packer;
    void *remoteBuf = nullptr; // receiver's unpack buffer, mapped into this process
    int remoteDev = -1;
    cudaEvent_t done; // interprocess event the receiver waits on
    bool awaitingAck = false; // the receiver has not yet unpacked the previous image
  };
  struct ColoRecv {
    int dst, srcRank, tagBase;
    std::unique_ptr<DeviceUnpacker> unpacker;
    cudaEvent_t senderDone = nullptr; // opened from the sender's handle
    bool pending = false;
  };
  std::vector<ColoSend> coloSends_;
  std::vector<ColoRecv> coloRecvs_;

  struct PlanLine {
    Route route;
    int src;
    Where dst;
    Dim3 dir;
    uint64_t bytes;
  };
  std::vector<PlanLine> planLines_;

  Where where(const Dim3 &idx) const {
    Placement *p = dd_.placement_.get();
    return Where{idx, p->get_rank(idx), p->get_subdomain_id(idx), p->get_cuda(idx)};
  }

  Route choose(const Where &me, const Where &other) const {
    const bool sameRank = other.rank == dd_.rank_;
    const bool sameNode = sameRank || dd_.topo_->colocated(other.rank);
    const bool sameDev = me.dev == other.dev;
    const bool peerOk = sameDev || gpu_topo::peer(me.dev, other.dev);
    if (dd_.any_methods(MethodFlags::CudaKernel) && sameRank && sameDev) return Route::Kernel;
    /* device-initiated stores: any peer-reachable GPU of this node. Two PROCESSES sharing one GPU are excluded:
       they would need spin-waits between kernels on the same device, which the hardware does not co-schedule. */
    if (dd_.any_methods(MethodFlags::NvlinkDirect) && sameNode && peerOk && (sameRank || !sameDev)) return Route::Direct;
    if (dd_.any_methods(MethodFlags::CudaMemcpyPeer) && sameRank && peerOk) return Route::PeerCopy;
    if (dd_.any_methods(MethodFlags::CudaMpiColocated) && !sameRank && sameNode && peerOk) return Route::Colo;
    if (dd_.any_methods(MethodFlags::CudaMpi | MethodFlags::CudaAwareMpi)) return Route::Staged;
    LOG_FATAL("No method available to send required message " << me.idx << " -> " << other.idx);
    return Route::Staged;
  }

  /* does a message travel from idx in direction dir (periodic unless the boundary says otherwise) */
  bool crosses_open_boundary(const Dim3 &idx, const Dim3 &dir, const Dim3 &grid) const {
    const Dim3 n = idx + dir;
    const bool outside = n.x < 0 || n.y < 0 || n.z < 0 || n.x >= grid.x || n.y >= grid.y || n.z >= grid.z;
    if (!outside) return false;
    /* which axes leave the domain? consult the boundary for that (axis-only) direction vector */
    const Dim3 leave(n.x < 0 ? -1 : (n.x >= grid.x ? 1 : 0), n.y < 0 ? -1 : (n.y >= grid.y ? 1 : 0),
                     n.z < 0 ? -1 : (n.z >= grid.z ? 1 : 0));
    for (int a = 0; a < 3; ++a) {
      if (leave[a] == 0) continue;
      Dim3 axisDir(0, 0, 0);
      axisDir[a] = leave[a];
      if (!dd_.boundary_.dir(int(axisDir.x), int(axisDir.y), int(axisDir.z))) return true;
    }
    return false;
  }

  /* view of quantity qi of subdomain `w` as seen from device myDev */
  BoxView view_of(const Where &w, size_t qi, int myDev) {
    BoxView v;
    if (w.rank == dd_.rank_) {
      v = dd_.domains_[size_t(w.dom)].view(qi);
    } else {
      RemoteSlab &rs = map_remote(myDev, w);
      v = rs.shadow->view_at(qi, static_cast<char *>(rs.ptr));
    }
    v.acrossLink = w.dev != myDev; // lets the plan spread the NVLink chunks over the whole launch
    return v;
  }

  /* dense landing box for the halo on side `side` of quantity qi of subdomain w, as addressed from device myDev */
  BoxView mailbox_of(const Where &w, const Dim3 &side, size_t qi, const Dim3 &ext, int myDev) {
    const LocalDomain *ld;
    char *base;
    if (w.rank == dd_.rank_) {
      ld = &dd_.domains_[size_t(w.dom)];
      base = static_cast<char *>(ld->slab());
    } else {
      RemoteSlab &rs = map_remote(myDev, w);
      ld = rs.shadow.get();
      base = static_cast<char *>(rs.ptr);
    }
    BoxView v = BoxView::dense(base + ld->mailbox_offset(side, qi, 0), ext, ld->elem_size(qi));
    v.base[1] = base + ld->mailbox_offset(side, qi, 1);
    v.acrossLink = w.dev != myDev;
    return v;
  }

  /* rows this short are sent to another GPU through its mailbox instead of straight into its halo */
  static bool thin_for_link(const Dim3 &dir, const Dim3 &ext, size_t elemSize) { return dir.x != 0 && size_t(ext.x) * elemSize <= 64 && elemSize % 4 == 0; }

  unsigned *flags_of(const Where &w, int myDev) {
    if (w.rank == dd_.rank_) return dd_.domains_[size_t(w.dom)].arrive_flags();
    return reinterpret_cast<unsigned *>(map_remote(myDev, w).ptr);
  }

  RemoteSlab &map_remote(int myDev, const Where &w) {
    const auto key = std::make_tuple(myDev, w.rank, w.dom);
    auto it = remote_.find(key);
    if (it != remote_.end()) return it->second;
    RemoteSlab rs;
    {
      DevGuard g(myDev);
      const cudaIpcMemHandle_t &h = allHandles_[size_t(domOffsetOfRank_[size_t(w.rank)] + w.dom)];
      CUDA_RUNTIME(cudaIpcOpenMemHandle(&rs.ptr, h, cudaIpcMemLazyEnablePeerAccess));
    }
    Placement *p = dd_.placement_.get();
    rs.shadow.reset(new LocalDomain(p->subdomain_size(w.idx), p->subdomain_origin(w.idx), w.dev));
    rs.shadow->set_radius(dd_.radius_);
    rs.shadow->set_alignment(dd_.align_);
    for (size_t qi = 0; qi < dd_.elemSizes_.size(); ++qi) rs.shadow->add_data(dd_.elemSizes_[qi], dd_.names_[qi]);
    rs.shadow->compute_layout();
    return remote_.emplace(key, std::move(rs)).first->second;
  }

public:
  explicit Exchanger(DistributedDomain &dd) : dd_(dd), nd_(dd.domains_.size()), timeoutNs_(flag_timeout_ns()) {
    if (const char *e = std::getenv("STENCIL_LAUNCH_THREADS")) threads_ = std::atoi(e) != 0;
  }

  ~Exchanger() {
    for (auto &kv : remote_) {
      DevGuard g(std::get<0>(kv.first));
      cudaIpcCloseMemHandle(kv.second.ptr);
    }
    for (cudaEvent_t e : pushDone_) cudaEventDestroy(e);
    for (cudaEvent_t e : exDone_) cudaEventDestroy(e);
    for (auto &pc : peerCopies_) cudaEventDestroy(pc.copied);
    for (auto &s : stagedSends_) {
      cudaEventDestroy(s.staged);
      cudaFreeHost(s.host);
    }
    for (auto &r : stagedRecvs_) cudaFreeHost(r.host);
    for (auto &c : coloSends_) {
      cudaEventDestroy(c.done);
      if (c.remoteBuf) cudaIpcCloseMemHandle(c.remoteBuf);
    }
    for (auto &c : coloRecvs_)
      if (c.senderDone) cudaEventDestroy(c.senderDone);
    if (errWord_) cudaFreeHost(errWord_);
  }

  cudaStream_t stream(size_t i) const { return streams_.at(i); }

  NeighborInfo neighbor(size_t di, const Dim3 &dir, size_t qi) {
    NeighborInfo n;
    Placement *pl = dd_.placement_.get();
    const Dim3 grid = pl->dim();
    const Where me = where(pl->get_idx(dd_.rank_, int(di)));
    if (crosses_open_boundary(me.idx, dir, grid)) return n;
    const Where nb = where((me.idx + dir).wrapped(grid));
    n.exists = true;
    n.rank = nb.rank;
    n.dom = nb.dom;
    n.dev = nb.dev;
    n.idx = nb.idx;
    n.size = pl->subdomain_size(nb.idx);
    n.sameRank = nb.rank == dd_.rank_;
    n.sameDevice = nb.dev == me.dev;
    const bool sameNode = n.sameRank || dd_.topo_->colocated(nb.rank);
    n.reachable = sameNode && (n.sameDevice || gpu_topo::peer(me.dev, nb.dev)) && (n.sameRank || !n.sameDevice);
    if (n.reachable) {
      n.view = view_of(nb, qi, me.dev);
      n.control = flags_of(nb, me.dev);
      if (n.sameRank) {
        n.layout = &dd_.domains_[size_t(nb.dom)];
        n.slab = static_cast<char *>(n.layout->slab());
      } else {
        RemoteSlab &rs = map_remote(me.dev, nb);
        n.layout = rs.shadow.get();
        n.slab = static_cast<char *>(rs.ptr);
      }
    }
    return n;
  }
  unsigned epoch() const { return epoch_; }
  const std::vector<PlanLine> &plan_lines() const { return planLines_; }

  /* ---------------- planning (realize) ---------------- */
  void prepare() {
    STENCIL_RANGE("DistributedDomain::realize() prep transports");
    Placement *pl = dd_.placement_.get();
    const Dim3 grid = pl->dim();
    stencil::Comm &comm = *dd_.comm_;

    CUDA_RUNTIME(cudaHostAlloc(&errWord_, sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable));
    *errWord_ = 0;

    if (const char *e = std::getenv("STENCIL_EXCHANGE_TRACE")) trace_ = std::atoi(e);
    for (size_t di = 0; di < nd_; ++di) {
      streams_.emplace_back(dd_.domains_[di].gpu(), RcStream::Priority::HIGH);
      DevGuard g(dd_.domains_[di].gpu());
      cudaEvent_t e;
      CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      pushDone_.push_back(e);
      cudaEvent_t d;
      CUDA_RUNTIME(cudaEventCreateWithFlags(&d, cudaEventDisableTiming));
      exDone_.push_back(d);
      after_.push_back(nullptr);
      if (trace_) {
        cudaEvent_t a, b, c;
        CUDA_RUNTIME(cudaEventCreate(&a));
        CUDA_RUNTIME(cudaEventCreate(&b));
        CUDA_RUNTIME(cudaEventCreate(&c));
        trA_.push_back(a);
        trB_.push_back(b);
        trC_.push_back(c);
      }
    }

    /* publish every rank's slab handles (also a barrier: all slabs exist and their flags are zeroed after this) */
    {
      std::vector<int> counts(static_cast<size_t>(comm.size()));
      const int mine = int(nd_);
      comm.allgather(&mine, counts.data(), sizeof(int));
      domOffsetOfRank_.assign(size_t(comm.size()), 0);
      int total = 0;
      std::vector<size_t> bytes(static_cast<size_t>(comm.size()));
      for (int r = 0; r < comm.size(); ++r) {
        domOffsetOfRank_[size_t(r)] = total;
        total += counts[size_t(r)];
        bytes[size_t(r)] = size_t(counts[size_t(r)]) * sizeof(cudaIpcMemHandle_t);
      }
      std::vector<cudaIpcMemHandle_t> mineH(nd_);
      if (comm.size() > 1)
        for (size_t di = 0; di < nd_; ++di) {
          DevGuard g(dd_.domains_[di].gpu());
          CUDA_RUNTIME(cudaIpcGetMemHandle(&mineH[di], dd_.domains_[di].slab()));
        }
      allHandles_.resize(size_t(total));
      comm.allgatherv(mineH.data(), allHandles_.data(), bytes);
    }

    push_.resize(nd_);
    scatter_.resize(nd_);
    arriveSig_.assign(nd_, SignalSet{0, {}});
    arriveWait_.assign(nd_, WaitSet{0, {}});
    readySig_.assign(nd_, SignalSet{0, {}});
    readyWait_.assign(nd_, WaitSet{0, {}});
    sameDevDeps_.assign(nd_, std::vector<int>());

    /* outboxes keyed by destination for the buffered transports */
    typedef std::pair<int, int> RD; // (rank, dom)
    std::vector<std::map<RD, std::vector<Message>>> peerOut(nd_), stagedOut(nd_), stagedIn(nd_), coloOut(nd_), coloIn(nd_);

    const Radius &radius = dd_.radius_;
    const size_t nq = dd_.elemSizes_.size();

    for (size_t di = 0; di < nd_; ++di) {
      LocalDomain &dom = dd_.domains_[di];
      const Where me = where(pl->get_idx(dd_.rank_, int(di)));
      std::set<int> depSet;
      for (int z = -1; z <= 1; ++z)
        for (int y = -1; y <= 1; ++y)
          for (int x = -1; x <= 1; ++x) {
            const Dim3 dir(x, y, z);
            if (dir == Dim3(0, 0, 0)) continue;
            const bool iSend = radius.dir(dir * -1) != 0 && !crosses_open_boundary(me.idx, dir, grid);
            /* the neighbour on side `dir` sends to me (travelling -dir) iff my halo on side dir exists */
            const bool iRecv = radius.dir(dir) != 0 && !crosses_open_boundary(me.idx, dir, grid);
            if (!iSend && !iRecv) continue;
            const Where nb = where((me.idx + dir).wrapped(grid));
            const Route route = choose(me, nb);

            if (iSend && dir.y == 0 && dir.z == 0) {
              ++xFaceMsgs_;
              if ((route == Route::Kernel || route == Route::Direct) && nb.dev == me.dev) ++xFaceLocal_;
            }
            if (iSend) {
              uint64_t bytes = 0;
              for (size_t qi = 0; qi < nq; ++qi) bytes += uint64_t(dom.halo_bytes(dir * -1, int64_t(qi)));
              planLines_.push_back(PlanLine{route, int(di), nb, dir, bytes});
              const Message msg(dir, int(di), nb.dom);
              switch (route) {
              case Route::Kernel:
              case Route::Direct: {
                /* my compute cells on side dir -> the neighbour's halo on its side -dir */
                const Dim3 ext = dom.halo_extent(dir * -1);
                const Dim3 srcPos = dom.halo_pos(dir, false);
                const Dim3 nbSize = pl->subdomain_size(nb.idx);
                const Dim3 dstPos = stencil::halo_pos(dir * -1, nbSize, radius, true);
                /* diagnostics only (results are wrong with it): STENCIL_EXCHANGE_DEBUG_DIRS is a mask of the message
                   classes to keep in the push plan — 1 x faces, 2 y faces, 4 z faces, 8 edges, 16 corners */
                if (const char *e = std::getenv("STENCIL_EXCHANGE_DEBUG_DIRS")) {
                  const int nz = (dir.x != 0) + (dir.y != 0) + (dir.z != 0);
                  const int cls = nz == 3 ? 16 : (nz == 2 ? 8 : (dir.x ? 1 : (dir.y ? 2 : 4)));
                  if (!(std::atoi(e) & cls)) break;
                }
                for (size_t qi = 0; qi < nq; ++qi) {
                  if (nb.dev != me.dev && thin_for_link(dir, ext, dd_.elemSizes_[qi]))
                    push_[di].add(dom.view(qi), srcPos, mailbox_of(nb, dir * -1, qi, ext, me.dev), Dim3(0, 0, 0), ext);
                  else if (nb.dev == me.dev && dir.y == 0 && dir.z == 0)
                    push_[di].add_optional(dom.view(qi), srcPos, view_of(nb, qi, me.dev), dstPos, ext);
                  else
                    push_[di].add(dom.view(qi), srcPos, view_of(nb, qi, me.dev), dstPos, ext);
                }
                break;
              }
              case Route::PeerCopy: peerOut[di][RD(nb.rank, nb.dom)].push_back(msg); break;
              case Route::Colo: coloOut[di][RD(nb.rank, nb.dom)].push_back(msg); break;
              case Route::Staged: stagedOut[di][RD(nb.rank, nb.dom)].push_back(msg); break;
              }
            }
            if (iRecv) {
              /* message from nb travelling in direction -dir */
              const Message msg(dir * -1, nb.dom, int(di));
              if ((route == Route::Direct || route == Route::Kernel) && nb.dev != me.dev) {
                /* the sender used my mailbox for the quantities whose rows are thin: scatter those into the halo */
                const Dim3 ext = dom.halo_extent(dir);
                for (size_t qi = 0; qi < nq; ++qi)
                  if (thin_for_link(dir, ext, dd_.elemSizes_[qi]))
                    scatter_[di].add(mailbox_of(me, dir, qi, ext, me.dev), Dim3(0, 0, 0), dom.view(qi), dom.halo_pos(dir, true), ext);
              }
              if (route == Route::Staged) stagedIn[di][RD(nb.rank, nb.dom)].push_back(msg);
              if (route == Route::Colo) coloIn[di][RD(nb.rank, nb.dom)].push_back(msg);
            }

            /* synchronisation edges of the direct path: symmetric (send OR receive) so that nobody overwrites a
               halo its owner may still be reading */
            if (route == Route::Kernel || route == Route::Direct) {
              const bool sameRank = nb.rank == dd_.rank_;
              if (sameRank && nb.dev == me.dev) {
                if (nb.dom != int(di)) depSet.insert(nb.dom);
              } else {
                const int slotMine = DirectionMap<int>::slot(dir);        // they raise this one in MY flags
                const int slotTheirs = DirectionMap<int>::slot(dir * -1); // I raise this one in THEIR flags
                unsigned *theirs = flags_of(nb, me.dev);
                arriveSig_[di].flag[arriveSig_[di].n++] = theirs + slotTheirs;
                arriveWait_[di].flag[arriveWait_[di].n++] = dom.arrive_flags() + slotMine;
                if (!sameRank) {
                  readySig_[di].flag[readySig_[di].n++] = theirs + 32 + slotTheirs;
                  readyWait_[di].flag[readyWait_[di].n++] = dom.arrive_flags() + 32 + slotMine;
                }
              }
            }
          }
      sameDevDeps_[di].assign(depSet.begin(), depSet.end());
      push_[di].finalize(dom.gpu());
      scatter_[di].finalize(dom.gpu());
    }

    /* peer copies: one packer/unpacker pair per (src, dst) local pair */
    for (size_t di = 0; di < nd_; ++di)
      for (auto &kv : peerOut[di]) {
        PeerCopy pc;
        pc.src = int(di);
        pc.dst = kv.first.second;
        pc.packer.reset(new DevicePacker(streams_[di]));
        pc.packer->prepare(&dd_.domains_[di], kv.second);
        pc.unpacker.reset(new DeviceUnpacker(streams_[size_t(pc.dst)]));
        pc.unpacker->prepare(&dd_.domains_[size_t(pc.dst)], kv.second);
        if (pc.packer->size() != pc.unpacker->size()) LOG_FATAL("peer copy size mismatch");
        DevGuard g(dd_.domains_[di].gpu());
        CUDA_RUNTIME(cudaEventCreateWithFlags(&pc.copied, cudaEventDisableTiming));
        peerCopies_.push_back(std::move(pc));
      }

    /* staged: tag = (src dom, dst dom); FIFO per (rank pair, tag) keeps iterations apart */
    deviceMsgs_ = dd_.any_methods(MethodFlags::CudaAwareMpi) && dd_.comm_->device_buffers();
    if (dd_.any_methods(MethodFlags::CudaAwareMpi) && !deviceMsgs_ && dd_.rank_ == 0)
      LOG_WARN("CudaAwareMpi requested but communicator '" << dd_.comm_->name() << "' does not take device pointers: staging through pinned memory");
    for (size_t di = 0; di < nd_; ++di) {
      for (auto &kv : stagedOut[di]) {
        StagedSend s;
        s.src = int(di);
        s.dstRank = kv.first.first;
        s.tag = staged_tag(int(di), kv.first.second);
        s.packer.reset(new DevicePacker(streams_[di]));
        s.packer->prepare(&dd_.domains_[di], kv.second);
        DevGuard g(dd_.domains_[di].gpu());
        if (!deviceMsgs_) CUDA_RUNTIME(cudaHostAlloc(&s.host, size_t(s.packer->size()), cudaHostAllocDefault));
        CUDA_RUNTIME(cudaEventCreateWithFlags(&s.staged, cudaEventDisableTiming));
        stagedSends_.push_back(std::move(s));
      }
      for (auto &kv : stagedIn[di]) {
        StagedRecv r;
        r.dst = int(di);
        r.srcRank = kv.first.first;
        r.tag = staged_tag(kv.first.second, int(di));
        r.unpacker.reset(new DeviceUnpacker(streams_[di]));
        r.unpacker->prepare(&dd_.domains_[di], kv.second);
        DevGuard g(dd_.domains_[di].gpu());
        if (!deviceMsgs_) CUDA_RUNTIME(cudaHostAlloc(&r.host, size_t(r.unpacker->size()), cudaHostAllocDefault));
        stagedRecvs_.push_back(std::move(r));
      }
    }

    /* colocated: receivers export their unpack buffer + device, senders export an interprocess event.
       Two-phase (everyone posts, then everyone collects) so no rank blocks another. */
    for (size_t di = 0; di < nd_; ++di) {
      for (auto &kv : coloIn[di]) {
        ColoRecv r;
        r.dst = int(di);
        r.srcRank = kv.first.first;
        r.tagBase = ((kv.first.second & 0xFF) << 8) | (int(di) & 0xFF);
        r.unpacker.reset(new DeviceUnpacker(streams_[di]));
        r.unpacker->prepare(&dd_.domains_[di], kv.second);
        DevGuard g(dd_.domains_[di].gpu());
        cudaIpcMemHandle_t mh;
        CUDA_RUNTIME(cudaIpcGetMemHandle(&mh, r.unpacker->data()));
        const int dev = dd_.domains_[di].gpu();
        comm.send(r.srcRank, make_tag<MsgKind::ColocatedMem>(r.tagBase), &mh, sizeof(mh));
        comm.send(r.srcRank, make_tag<MsgKind::ColocatedDev>(r.tagBase), &dev, sizeof(dev));
        coloRecvs_.push_back(std::move(r));
      }
      for (auto &kv : coloOut[di]) {
        ColoSend s;
        s.src = int(di);
        s.dstRank = kv.first.first;
        s.tagBase = ((int(di) & 0xFF) << 8) | (kv.first.second & 0xFF);
        s.packer.reset(new DevicePacker(streams_[di]));
        s.packer->prepare(&dd_.domains_[di], kv.second);
        DevGuard g(dd_.domains_[di].gpu());
        CUDA_RUNTIME(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming | cudaEventInterprocess));
        cudaIpcEventHandle_t eh;
        CUDA_RUNTIME(cudaIpcGetEventHandle(&eh, s.done));
        comm.send(s.dstRank, make_tag<MsgKind::ColocatedEvt>(s.tagBase), &eh, sizeof(eh));
        coloSends_.push_back(std::move(s));
      }
    }
    for (ColoSend &s : coloSends_) {
      cudaIpcMemHandle_t mh;
      comm.recv(s.dstRank, make_tag<MsgKind::ColocatedMem>(s.tagBase), &mh, sizeof(mh));
      comm.recv(s.dstRank, make_tag<MsgKind::ColocatedDev>(s.tagBase), &s.remoteDev, sizeof(int));
      DevGuard g(dd_.domains_[size_t(s.src)].gpu());
      CUDA_RUNTIME(cudaIpcOpenMemHandle(&s.remoteBuf, mh, cudaIpcMemLazyEnablePeerAccess));
    }
    for (ColoRecv &r : coloRecvs_) {
      cudaIpcEventHandle_t eh;
      comm.recv(r.srcRank, make_tag<MsgKind::ColocatedEvt>(r.tagBase), &eh, sizeof(eh));
      DevGuard g(dd_.domains_[size_t(r.dst)].gpu());
      CUDA_RUNTIME(cudaIpcOpenEventHandle(&r.senderDone, eh));
    }

    for (size_t di = 0; di < nd_; ++di) {
      DevGuard g(dd_.domains_[di].gpu());
      CUDA_RUNTIME(cudaDeviceSynchronize());
    }
    comm.barrier();
  }

  /* ---------------- the per-iteration exchange ---------------- */
  void exchange() {
    STENCIL_RANGE("DD::exchange()");
    enqueue();
    finish();
  }

  bool device_only() const { return stagedSends_.empty() && stagedRecvs_.empty() && coloSends_.empty() && coloRecvs_.empty(); }
  bool in_flight() const { return inFlight_; }
  void skip_local_x_faces(bool on) { skipXLocal_ = on; }
  bool local_x_faces_only() const { // every +-x face message of every local subdomain is optional
    return xFaceMsgs_ > 0 && xFaceMsgs_ == xFaceLocal_;
  }
  cudaEvent_t done_event(size_t di) const { return exDone_[di]; }
  void set_after(size_t di, cudaEvent_t e) { after_[di] = e; }

  /* Everything exchange() puts on the GPUs, without waiting for any of it. */
  void enqueue() {
    STENCIL_RANGE("DD::exchange: enqueue");
    if (inFlight_) {
      if (!device_only()) LOG_FATAL("exchange_async(): the previous exchange uses host-mediated transports and was not completed with exchange_wait()");
      check_error_word();
    }
    inFlight_ = true;
    ++epoch_;
    stencil::Comm &comm = *dd_.comm_;
    std::vector<LocalDomain> &doms = dd_.domains_;

    /* work the caller declared with exchange_after(): producers of the data to send, and consumers of the halos about to
       be overwritten. Any local subdomain may receive from any other, so every exchange stream waits for all of them. */
    for_each_dom([&](size_t di) {
      DevGuard g(doms[di].gpu());
      for (size_t dj = 0; dj < nd_; ++dj)
        if (after_[dj]) CUDA_RUNTIME(cudaStreamWaitEvent(streams_[di], after_[dj], 0));
      /* 0. tell cross-process senders that my halos are free for this epoch (only other processes can be ahead of me) */
      if (readySig_[di].n) launch_signal_flags(readySig_[di], epoch_, streams_[di]);
    });

    /* 1. longest paths first: staged sends (pack + D2H) */
    {
    STENCIL_RANGE("DD::exchange: remote send d2h");
    for (StagedSend &s : stagedSends_) {
      DevGuard g(doms[size_t(s.src)].gpu());
      s.packer->pack();
      if (!deviceMsgs_)
        CUDA_RUNTIME(cudaMemcpyAsync(s.host, s.packer->data(), size_t(s.packer->size()), cudaMemcpyDeviceToHost, streams_[size_t(s.src)]));
      CUDA_RUNTIME(cudaEventRecord(s.staged, streams_[size_t(s.src)]));
      s.pending = true;
    }
    }
    for (StagedRecv &r : stagedRecvs_) r.pending = true;

    /* 2. colocated sends: pack, copy into the receiver's buffer, record the shared event, notify */
    for (ColoSend &s : coloSends_) {
      STENCIL_RANGE("DD::exchange: colo send");
      DevGuard g(doms[size_t(s.src)].gpu());
      cudaStream_t st = streams_[size_t(s.src)];
      if (s.awaitingAck) { // do not overwrite an image the receiver has not unpacked yet
        char ack;
        comm.recv(s.dstRank, ack_tag(s.tagBase), &ack, 1);
      }
      s.awaitingAck = true;
      s.packer->pack();
      CUDA_RUNTIME(cudaMemcpyPeerAsync(s.remoteBuf, s.remoteDev, s.packer->data(), doms[size_t(s.src)].gpu(), size_t(s.packer->size()), st));
      CUDA_RUNTIME(cudaEventRecord(s.done, st));
      const char token = 1;
      comm.send(s.dstRank, make_tag<MsgKind::ColocatedNotify>(s.tagBase), &token, 1);
    }
    for (ColoRecv &r : coloRecvs_) r.pending = true;

    /* 3. peer copies, phase A: pack + copy + record on the source streams */
    for (PeerCopy &pc : peerCopies_) {
      STENCIL_RANGE("DD::exchange: peer copy send");
      DevGuard g(doms[size_t(pc.src)].gpu());
      cudaStream_t st = streams_[size_t(pc.src)];
      pc.packer->pack();
      CUDA_RUNTIME(cudaMemcpyPeerAsync(pc.unpacker->data(), doms[size_t(pc.dst)].gpu(), pc.packer->data(), doms[size_t(pc.src)].gpu(),
                                       size_t(pc.packer->size()), st));
      CUDA_RUNTIME(cudaEventRecord(pc.copied, st));
    }

    /* 4. direct pushes: wait for cross-process credits, store into the neighbours' halos, raise their flags */
    for_each_dom([&](size_t di) {
      if (push_[di].empty() && arriveSig_[di].n == 0) return;
      STENCIL_RANGE("DD::exchange: direct push");
      DevGuard g(doms[di].gpu());
      launch_wait_flags(readyWait_[di], epoch_, errWord_, timeoutNs_, streams_[di]);
      if (trace_) CUDA_RUNTIME(cudaEventRecord(trA_[di], streams_[di]));
      push_[di].launch_and_signal(doms[di].parity(), arriveSig_[di], epoch_, streams_[di], !skipXLocal_);
      if (trace_) CUDA_RUNTIME(cudaEventRecord(trB_[di], streams_[di]));
      CUDA_RUNTIME(cudaEventRecord(pushDone_[di], streams_[di]));
    });

    /* 5. phase B: everything that waits on another stream is enqueued only now, so no stream can wait on work
          that has not been submitted */
    for (PeerCopy &pc : peerCopies_) {
      DevGuard g(doms[size_t(pc.dst)].gpu());
      CUDA_RUNTIME(cudaStreamWaitEvent(streams_[size_t(pc.dst)], pc.copied, 0));
      pc.unpacker->unpack();
    }
    for_each_dom([&](size_t di) {
      DevGuard g(doms[di].gpu());
      for (int dep : sameDevDeps_[di]) CUDA_RUNTIME(cudaStreamWaitEvent(streams_[di], pushDone_[size_t(dep)], 0));
      launch_wait_flags(arriveWait_[di], epoch_, errWord_, timeoutNs_, streams_[di]);
      scatter_[di].launch(doms[di].parity(), streams_[di]);
      if (trace_) CUDA_RUNTIME(cudaEventRecord(trC_[di], streams_[di]));
    });

    if (device_only())
      for (size_t di = 0; di < nd_; ++di) {
        DevGuard g(doms[di].gpu());
        CUDA_RUNTIME(cudaEventRecord(exDone_[di], streams_[di]));
      }
    for (size_t di = 0; di < nd_; ++di) after_[di] = nullptr; // one-shot
  }

  /* Host side of the exchange: progress of the host-mediated transports, then completion of the streams. */
  void finish() {
    if (!inFlight_) return;
    stencil::Comm &comm = *dd_.comm_;
    std::vector<LocalDomain> &doms = dd_.domains_;
    /* 6. host progress loop for the host-mediated transports */
    size_t open = stagedSends_.size() + stagedRecvs_.size() + coloRecvs_.size();
    STENCIL_RANGE("DD::exchange: poll");
    const double t0 = stencil::Comm::wtime();
    while (open) {
      for (StagedSend &s : stagedSends_) {
        if (!s.pending) continue;
        const cudaError_t q = cudaEventQuery(s.staged);
        if (q == cudaSuccess) {
          comm.send(s.dstRank, s.tag, deviceMsgs_ ? s.packer->data() : s.host, size_t(s.packer->size()));
          s.pending = false;
          --open;
        } else if (q != cudaErrorNotReady) {
          CUDA_RUNTIME(q);
        }
      }
      for (StagedRecv &r : stagedRecvs_) {
        if (!r.pending) continue;
        if (comm.try_recv(r.srcRank, r.tag, deviceMsgs_ ? r.unpacker->data() : r.host, size_t(r.unpacker->size()))) {
          DevGuard g(doms[size_t(r.dst)].gpu());
          if (!deviceMsgs_)
            CUDA_RUNTIME(cudaMemcpyAsync(r.unpacker->data(), r.host, size_t(r.unpacker->size()), cudaMemcpyHostToDevice, streams_[size_t(r.dst)]));
          r.unpacker->unpack();
          r.pending = false;
          --open;
        }
      }
      for (ColoRecv &r : coloRecvs_) {
        if (!r.pending) continue;
        char token;
        if (comm.try_recv(r.srcRank, make_tag<MsgKind::ColocatedNotify>(r.tagBase), &token, 1)) {
          DevGuard g(doms[size_t(r.dst)].gpu());
          CUDA_RUNTIME(cudaStreamWaitEvent(streams_[size_t(r.dst)], r.senderDone, 0));
          r.unpacker->unpack();
          r.pending = false;
          --open;
        }
      }
      if (open && stencil::Comm::wtime() - t0 > 120.0) LOG_FATAL("exchange(): host-mediated transports made no progress for 120 s");
    }

    /* every receive of mine is in: in-place (device-pointer) sends may now be completed without two ranks waiting for
       each other, and must be before the packers are filled again */
    if (deviceMsgs_) comm.flush_sends();

    /* 7. completion */
    STENCIL_RANGE("DD::exchange: wait");
    for (size_t di = 0; di < nd_; ++di) {
      DevGuard g(doms[di].gpu());
      CUDA_RUNTIME(cudaStreamSynchronize(streams_[di]));
    }
    if (trace_ > 0) {
      --trace_;
      for (size_t di = 0; di < nd_; ++di) {
        if (push_[di].empty()) continue;
        float push = 0, wait = 0;
        cudaEventElapsedTime(&push, trA_[di], trB_[di]);
        cudaEventElapsedTime(&wait, trB_[di], trC_[di]);
        std::fprintf(stderr, "exchange %u rank %d dom %zu: push %.1f us (%zu B, %zu segs, %zu chunks), flag wait %.1f us\n", epoch_, dd_.rank_, di,
                     push * 1e3, push_[di].bytes(), push_[di].num_segments(), push_[di].num_chunks(), wait * 1e3);
      }
    }
    for (ColoRecv &r : coloRecvs_) {
      const char ack = 1;
      comm.send(r.srcRank, ack_tag(r.tagBase), &ack, 1);
    }
    inFlight_ = false;
    check_error_word();
  }

  void check_error_word() const {
    if (*errWord_ != 0) {
      LOG_FATAL("exchange(): timed out waiting for a neighbour's flag (code " << *errWord_ << ", epoch " << epoch_
                                                                             << "); a peer process is gone or stuck");
    }
  }

  /* ---------------- plan dump + byte counters ---------------- */
  void summarize() {
    STENCIL_RANGE("allreduce communication stats");
    std::vector<LocalDomain> &doms = dd_.domains_;
    Placement *pl = dd_.placement_.get();
    uint64_t staged = 0, colo = 0, peer = 0, kernel = 0, direct = 0;
    for (const PlanLine &l : planLines_) {
      switch (l.route) {
      case Route::Kernel: kernel += l.bytes; break;
      case Route::Direct: direct += l.bytes; break;
      case Route::PeerCopy: peer += l.bytes; break;
      case Route::Colo: colo += l.bytes; break;
      case Route::Staged: staged += l.bytes; break;
      }
    }
    if (dd_.planFile_) {
      std::ofstream f("plan_" + std::to_string(dd_.rank_) + ".txt");
      f << "rank=" << dd_.rank_ << "\n\n";
      f << "== quantities == \n";
      for (size_t qi = 0; qi < dd_.elemSizes_.size(); ++qi)
        f << qi << ":" << (dd_.names_[qi].empty() ? "data" + std::to_string(qi) : dd_.names_[qi]) << ":" << dd_.elemSizes_[qi] << "B\n";
      f << "domains\n";
      for (size_t di = 0; di < nd_; ++di)
        f << di << ":cuda" << doms[di].gpu() << ":" << pl->get_idx(dd_.rank_, int(di)) << " sz=" << doms[di].size() << "\n";
      f << "\n";
      const struct {
        Route r;
        const char *title;
      } sections[] = {{Route::Kernel, "== peerAccess =="},
                      {Route::Direct, "== nvlinkDirect =="},
                      {Route::PeerCopy, "== peerCopy =="},
                      {Route::Colo, "== colo =="},
                      {Route::Staged, "== remote =="}};
      for (const auto &sec : sections) {
        f << sec.title << "\n";
        for (const PlanLine &l : planLines_) {
          if (l.route != sec.r) continue;
          if (sec.r == Route::Colo || sec.r == Route::Staged)
            f << route_name(sec.r) << " to dstIdx=" << l.dst.idx << " dir=" << l.dir << " (" << l.src << "->" << l.dst.dom << ") "
              << l.bytes << "B\n";
          else
            f << l.src << "->" << l.dst.dom << " " << l.dir << " " << l.bytes << "B\n";
        }
        f << "\n";
      }
    }
    stencil::Comm &comm = *dd_.comm_;
    /* deviceMsgs_ is a property of the communicator, so it is the same on every rank */
    dd_.numBytesCudaMpi_ = comm.allreduce_sum(deviceMsgs_ ? 0 : staged);
    dd_.numBytesCudaAwareMpi_ = comm.allreduce_sum(deviceMsgs_ ? staged : 0);
    dd_.numBytesCudaMpiColocated_ = comm.allreduce_sum(colo);
    dd_.numBytesCudaMemcpyPeer_ = comm.allreduce_sum(peer);
    dd_.numBytesCudaKernel_ = comm.allreduce_sum(kernel);
    dd_.numBytesNvlinkDirect_ = comm.allreduce_sum(direct);
    if (dd_.rank_ == 0) {
      LOG_INFO(dd_.numBytesCudaMpi_ << "B CudaMpi / exchange");
      LOG_INFO(dd_.numBytesCudaAwareMpi_ << "B CudaAwareMpi / exchange");
      LOG_INFO(dd_.numBytesCudaMpiColocated_ << "B CudaMpiColocated / exchange");
      LOG_INFO(dd_.numBytesCudaMemcpyPeer_ << "B CudaMemcpyPeer / exchange");
      LOG_INFO(dd_.numBytesCudaKernel_ << "B CudaKernel / exchange");
      LOG_INFO(dd_.numBytesNvlinkDirect_ << "B NvlinkDirect / exchange");
    }
  }
};

} // namespace stencil

/* ================================================================================================================ */

DistributedDomain::DistributedDomain(size_t x, size_t y, size_t z) : comm_(&stencil::world()) { init(x, y, z); }
DistributedDomain::DistributedDomain(size_t x, size_t y, size_t z, stencil::Comm &comm) : comm_(&comm) { init(x, y, z); }
DistributedDomain::~DistributedDomain() {
  ex_.reset(); // transports reference the domains: tear them down first
  domains_.clear();
}

/* Setup timers (reference: STENCIL_SETUP_STATS, default ON, stencil.cu:31-52 there): max over ranks of each phase's
   wall time, valid on every rank. Built with -DSETUP_STATS=OFF the reductions (one small collective per phase) are
   skipped and the fields hold this rank's own times. */
static double setup_phase_seconds(stencil::Comm &comm, double local) {
#if !defined(STENCIL_SETUP_STATS) || STENCIL_SETUP_STATS == 1
  return comm.allreduce_max(local);
#else
  (void)comm;
  return local;
#endif
}

void DistributedDomain::init(size_t x, size_t y, size_t z) {
  size_ = Dim3(int64_t(x), int64_t(y), int64_t(z));
  rank_ = comm_->rank();
  worldSize_ = comm_->size();

  double t = stencil::Comm::wtime();
  topo_.reset(new MpiTopology(*comm_));
  timeMpiTopo_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  int deviceCount = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&deviceCount));
  if (deviceCount <= 0) LOG_FATAL("no CUDA devices");
  LOG_DEBUG("rank " << rank_ << ": " << deviceCount << " CUDA devices, colocated rank " << topo_->colocated_rank() << "/"
                    << topo_->colocated_size());

  /* default device choice: spread the node's GPUs over the node's ranks; oversubscribe round-robin if needed */
  const int cr = topo_->colocated_rank(), cs = topo_->colocated_size();
  if (cs <= deviceCount) {
    for (int id = 0; id < deviceCount; ++id)
      if (id % cs == cr) gpus_.push_back(id);
  } else {
    gpus_.push_back(cr % deviceCount);
  }
}

void DistributedDomain::set_gpus(const std::vector<int> &cudaIds) { gpus_ = cudaIds; }

Dim3 DistributedDomain::get_subdomain_idx(int64_t i) const { return placement_->get_idx(rank_, int(i)); }
Dim3 DistributedDomain::subdomain_grid() const { return placement_->dim(); }
cudaStream_t DistributedDomain::exchange_stream(size_t i) const { return ex_->stream(i); }
unsigned DistributedDomain::epoch() const noexcept { return ex_ ? ex_->epoch() : 0; }

stencil::NeighborInfo DistributedDomain::neighbor(size_t di, const Dim3 &dir, size_t qi) { return ex_->neighbor(di, dir, qi); }

uint64_t DistributedDomain::exchange_bytes_for_method(const MethodFlags &m) const {
  uint64_t r = 0;
  if (m && MethodFlags::CudaMpi) r += numBytesCudaMpi_;
  if (m && MethodFlags::CudaAwareMpi) r += numBytesCudaAwareMpi_;
  if (m && MethodFlags::CudaMpiColocated) r += numBytesCudaMpiColocated_;
  if (m && MethodFlags::CudaMemcpyPeer) r += numBytesCudaMemcpyPeer_;
  if (m && MethodFlags::CudaKernel) r += numBytesCudaKernel_;
  if (m && MethodFlags::NvlinkDirect) r += numBytesNvlinkDirect_;
  return r;
}

void DistributedDomain::realize() {
  const double tRealize = stencil::Comm::wtime();

  /* node-wide list of contributed devices, then peer access between mine and all of them */
  double t = stencil::Comm::wtime();
  std::vector<int> nodeCuda;
  {
    std::vector<int> counts(static_cast<size_t>(worldSize_));
    const int mine = int(gpus_.size());
    comm_->allgather(&mine, counts.data(), sizeof(int));
    std::vector<size_t> bytes(static_cast<size_t>(worldSize_));
    int total = 0;
    for (int r = 0; r < worldSize_; ++r) {
      bytes[size_t(r)] = size_t(counts[size_t(r)]) * sizeof(int);
      total += counts[size_t(r)];
    }
    std::vector<int> all(static_cast<size_t>(total));
    comm_->allgatherv(gpus_.data(), all.data(), bytes);
    int off = 0;
    for (int r = 0; r < worldSize_; ++r) {
      if (topo_->colocated(r))
        for (int i = 0; i < counts[size_t(r)]; ++i) nodeCuda.push_back(all[size_t(off + i)]);
      off += counts[size_t(r)];
    }
  }
  timeNodeGpus_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  t = stencil::Comm::wtime();
  {
    STENCIL_RANGE("peer_en");
    /* peer access between the GPUs of THIS process (both directions: the loop visits every ordered pair); GPUs of
       other processes are only noted as reachable — their slabs arrive as IPC mappings that enable access lazily */
    for (int src : gpus_)
      for (int dst : nodeCuda) {
        if (std::find(gpus_.begin(), gpus_.end(), dst) != gpus_.end())
          gpu_topo::enable_peer(src, dst);
        else
          gpu_topo::note_peer(src, dst);
      }
  }
  timePeerEn_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  t = stencil::Comm::wtime();
  {
    STENCIL_RANGE("placement");
    if (strategy_ == PlacementStrategy::NodeAware)
      placement_.reset(new NodeAware(size_, *topo_, radius_, gpus_, bandwidthFn_));
    else
      placement_.reset(new Trivial(size_, *topo_, gpus_));
  }
  timePlacement_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  t = stencil::Comm::wtime();
  domains_.clear();
  domains_.reserve(gpus_.size());
  for (size_t di = 0; di < gpus_.size(); ++di) {
    const Dim3 idx = placement_->get_idx(rank_, int(di));
    const int cuda = placement_->get_cuda(idx);
    LocalDomain ld(placement_->subdomain_size(idx), placement_->subdomain_origin(idx), cuda);
    ld.set_radius(radius_);
    ld.set_alignment(align_);
    for (size_t qi = 0; qi < elemSizes_.size(); ++qi) ld.add_data(elemSizes_[qi], names_[qi]);
    if (stencil::log::runtime_level() >= 3)
      std::fprintf(stderr, "rank=%d gpu=%zu (cuda id=%d) => [%ld,%ld,%ld]\n", rank_, di, cuda, long(idx.x), long(idx.y), long(idx.z));
    domains_.push_back(std::move(ld));
  }
  for (LocalDomain &d : domains_) d.realize();
  timeCreate_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  t = stencil::Comm::wtime();
  STENCIL_RANGE("DistributedDomain::realize() plan messages");
  ex_.reset(new stencil::Exchanger(*this));
  ex_->prepare();
  ex_->summarize();
  timePlan_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - t);

  timeRealize_ += setup_phase_seconds(*comm_, stencil::Comm::wtime() - tRealize);
}

void DistributedDomain::swap() {
  double t = 0;
  if (exchangeStats_) {
    comm_->barrier();
    t = stencil::Comm::wtime();
  }
  {
    STENCIL_RANGE("swap");
    for (LocalDomain &d : domains_) d.swap();
  }
  if (exchangeStats_) timeSwap_ += comm_->allreduce_max(stencil::Comm::wtime() - t);
}

void DistributedDomain::exchange_async() { ex_->enqueue(); }
void DistributedDomain::exchange_wait() { ex_->finish(); }
bool DistributedDomain::exchange_device_only() const { return ex_->device_only(); }
cudaEvent_t DistributedDomain::exchange_done_event(size_t i) const {
  if (!ex_->device_only()) LOG_FATAL("exchange_done_event(): this plan has host-mediated transports; use exchange_wait()");
  return ex_->done_event(i);
}
void DistributedDomain::exchange_after(size_t i, cudaEvent_t e) { ex_->set_after(i, e); }
void DistributedDomain::exchange_skip_local_x_faces(bool on) { ex_->skip_local_x_faces(on); }
bool DistributedDomain::exchange_x_faces_all_local() const { return ex_->local_x_faces_only(); }

void DistributedDomain::exchange() {
  double t = 0;
  if (exchangeStats_) {
    comm_->barrier();
    t = stencil::Comm::wtime();
  }
  ex_->exchange();
  if (exchangeStats_) timeExchange_ += comm_->allreduce_max(stencil::Comm::wtime() - t);
}

double DistributedDomain::time_exchange(int iters, double *deviceSeconds) {
  std::vector<cudaEvent_t> t0(domains_.size()), t1(domains_.size());
  comm_->barrier();
  for (size_t di = 0; di < domains_.size(); ++di) {
    stencil::DevGuard g(domains_[di].gpu());
    CUDA_RUNTIME(cudaEventCreate(&t0[di]));
    CUDA_RUNTIME(cudaEventCreate(&t1[di]));
    CUDA_RUNTIME(cudaEventRecord(t0[di], exchange_stream(di)));
  }
  double elapsed = stencil::Comm::wtime();
  for (int i = 0; i < iters; ++i) ex_->exchange();
  elapsed = comm_->allreduce_max(stencil::Comm::wtime() - elapsed);
  float worst = 0;
  for (size_t di = 0; di < domains_.size(); ++di) {
    stencil::DevGuard g(domains_[di].gpu());
    CUDA_RUNTIME(cudaEventRecord(t1[di], exchange_stream(di)));
    CUDA_RUNTIME(cudaEventSynchronize(t1[di]));
    float ms = 0;
    CUDA_RUNTIME(cudaEventElapsedTime(&ms, t0[di], t1[di]));
    worst = std::max(worst, ms);
    cudaEventDestroy(t0[di]);
    cudaEventDestroy(t1[di]);
  }
  if (deviceSeconds) *deviceSeconds = comm_->allreduce_max(double(worst) * 1e-3);
  return elapsed;
}

std::vector<Rect3> DistributedDomain::get_interior() const {
  std::vector<Rect3> out;
  for (const LocalDomain &d : domains_) {
    const Rect3 com = d.get_compute_region();
    Rect3 in = com;
    /* a cell is interior if no stencil access leaves the compute region: shrink every side by the largest radius
       among the directions that point to that side */
    for (int z = -1; z <= 1; ++z)
      for (int y = -1; y <= 1; ++y)
        for (int x = -1; x <= 1; ++x) {
          const int c[3] = {x, y, z};
          if (!x && !y && !z) continue;
          const int64_t r = int64_t(radius_.dir(x, y, z));
          for (int a = 0; a < 3; ++a) {
            if (c[a] < 0) in.lo[a] = std::max(in.lo[a], com.lo[a] + r);
            if (c[a] > 0) in.hi[a] = std::min(in.hi[a], com.hi[a] - r);
          }
        }
    out.push_back(in);
  }
  return out;
}

std::vector<std::vector<Rect3>> DistributedDomain::get_exterior() const {
  std::vector<std::vector<Rect3>> out(domains_.size());
  const std::vector<Rect3> interiors = get_interior();
  for (size_t di = 0; di < domains_.size(); ++di) {
    const Rect3 &in = interiors[di];
    Rect3 rest = domains_[di].get_compute_region(); // shrinks as slabs are peeled off
    /* peel +x, +y, +z, -x, -y, -z: each slab spans what is left in the other axes, so slabs never overlap */
    for (int side = 0; side < 6; ++side) {
      const int a = side % 3;
      const bool positive = side < 3;
      Rect3 slab = rest;
      if (positive) {
        if (in.hi[a] == rest.hi[a]) continue;
        slab.lo[a] = in.hi[a];
        rest.hi[a] = in.hi[a];
      } else {
        if (in.lo[a] == rest.lo[a]) continue;
        slab.hi[a] = in.lo[a];
        rest.lo[a] = in.lo[a];
      }
      out[di].push_back(slab);
    }
  }
  return out;
}

void DistributedDomain::write_paraview(const std::string &prefix, bool zeroNaNs) {
  STENCIL_RANGE("write_paraview");
  for (size_t di = 0; di < domains_.size(); ++di) {
    const int64_t id = int64_t(rank_) * int64_t(domains_.size()) + int64_t(di);
    const std::string path = prefix + "_" + std::to_string(id) + ".txt";
    LocalDomain &d = domains_[di];
    std::vector<std::vector<unsigned char>> q;
    for (int64_t qi = 0; qi < d.num_data(); ++qi) q.push_back(d.interior_to_host(size_t(qi)));
    LOG_INFO("open " << path);
    FILE *f = std::fopen(path.c_str(), "w");
    if (!f) LOG_FATAL("cannot open " << path);
    std::fprintf(f, "Z,Y,X");
    for (int64_t qi = 0; qi < d.num_data(); ++qi) {
      const std::string &n = d.data_name(size_t(qi));
      std::fprintf(f, ",%s", n.empty() ? ("data" + std::to_string(qi)).c_str() : n.c_str());
    }
    std::fprintf(f, "\n");
    const Dim3 sz = d.size(), org = d.origin();
    for (int64_t lz = 0; lz < sz.z; ++lz)
      for (int64_t ly = 0; ly < sz.y; ++ly)
        for (int64_t lx = 0; lx < sz.x; ++lx) {
          const size_t i = size_t((lz * sz.y + ly) * sz.x + lx);
          std::fprintf(f, "%ld,%ld,%ld", long(org.z + lz), long(org.y + ly), long(org.x + lx));
          for (int64_t qi = 0; qi < d.num_data(); ++qi) {
            double v;
            if (d.elem_size(size_t(qi)) == 8)
              v = reinterpret_cast<const double *>(q[size_t(qi)].data())[i];
            else if (d.elem_size(size_t(qi)) == 4)
              v = double(reinterpret_cast<const float *>(q[size_t(qi)].data())[i]);
            else
              continue;
            if (zeroNaNs && std::isnan(v)) v = 0.0;
            std::fprintf(f, ",%f", v);
          }
          std::fprintf(f, "\n");
        }
    std::fclose(f);
  }
}
