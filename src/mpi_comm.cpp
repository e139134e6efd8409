/* MpiComm: stencil::Comm over a real MPI, for multi-node runs (mpirun / srun / jsrun launchers as in the
   reference's scripts/{hal,summit}). Compiled to a stub unless the build was configured with -DUSE_MPI=ON.

   Mapping onto the Comm contract:
     send       eager: the payload is copied into a heap buffer that is MPI_Isend'ed and reaped later, so send()
                never waits for the receiver (the reference posts MPI_Isend and polls, tx_cuda.cuh:646,599).
                With a CUDA-aware MPI and a device pointer the user buffer is sent in place and completed before
                returning (the packed buffer is reused by the next exchange).
     try_recv   MPI_Iprobe(src, tag) + MPI_Recv — the Irecv/Test pair of tx_cuda.cuh:768,748.
     collectives  the MPI ones (stencil.hpp:151-249, partition.hpp:390-835 of the reference).
   Tags: the library's tags fit in 15 bits plus a kind field (tx_common.hpp), below MPI_TAG_UB's guaranteed 32767
   after folding; see fold(). */
#include "stencil/comm.hpp"

#if defined(STENCIL_USE_MPI) && STENCIL_USE_MPI == 1

#include <mpi.h>

#include <cstdlib>
#include <cstring>
#include <list>
#include <vector>

#include "stencil/logging.hpp"

namespace stencil {
namespace {

class MpiComm : public Comm {
public:
  MpiComm() {
    int inited = 0;
    MPI_Initialized(&inited);
    if (!inited) {
      int provided = 0;
      MPI_Init_thread(nullptr, nullptr, MPI_THREAD_FUNNELED, &provided);
      owner_ = true;
    }
    MPI_Comm_dup(MPI_COMM_WORLD, &comm_);
    MPI_Comm_rank(comm_, &rank_);
    MPI_Comm_size(comm_, &size_);
    int flag = 0;
    void *ub = nullptr;
    MPI_Comm_get_attr(MPI_COMM_WORLD, MPI_TAG_UB, &ub, &flag);
    tagUb_ = (flag && ub) ? *static_cast<int *>(ub) : 32767;
  }
  ~MpiComm() override {
    for (Pending &p : pending_) MPI_Wait(&p.req, MPI_STATUS_IGNORE);
    pending_.clear();
    MPI_Comm_free(&comm_);
    if (owner_) {
      int fin = 0;
      MPI_Finalized(&fin);
      if (!fin) MPI_Finalize();
    }
  }

  int rank() const override { return rank_; }
  int size() const override { return size_; }
  const char *name() const override { return "mpi"; }
  bool device_buffers() const override {
#if defined(STENCIL_USE_CUDA_AWARE_MPI) && STENCIL_USE_CUDA_AWARE_MPI == 1
    return true;
#else
    return false;
#endif
  }

  void send(int dst, int tag, const void *buf, size_t bytes) override {
    reap();
    if (bytes > size_t(2147483647)) LOG_FATAL("MpiComm::send: message of " << bytes << " B exceeds INT_MAX");
    if (device_buffers()) {
      /* may be a device pointer: it cannot be copied with memcpy, so it is sent in place — and NOT completed here:
         device buffers usually go by rendezvous, and two ranks that both wait for their send before polling a receive
         would wait for each other for ever (the reference posts Isend and polls, tx_cuda.cuh:776-974). The request is
         reaped by later calls; flush_sends() completes it before the caller reuses the buffer. */
      pending_.emplace_back();
      pending_.back().inPlace = true;
      MPI_Isend(buf, int(bytes), MPI_BYTE, dst, fold(tag), comm_, &pending_.back().req);
      return;
    }
    pending_.emplace_back();
    Pending &p = pending_.back();
    p.data.assign(static_cast<const char *>(buf), static_cast<const char *>(buf) + bytes);
    MPI_Isend(p.data.data(), int(bytes), MPI_BYTE, dst, fold(tag), comm_, &p.req);
  }

  void flush_sends() override {
    for (;;) {
      reap();
      bool open = false;
      for (const Pending &p : pending_) open = open || p.inPlace;
      if (!open) return;
    }
  }

  bool try_recv(int src, int tag, void *buf, size_t bytes) override {
    reap();
    int flag = 0;
    MPI_Status st;
    MPI_Iprobe(src, fold(tag), comm_, &flag, &st);
    if (!flag) return false;
    int count = 0;
    MPI_Get_count(&st, MPI_BYTE, &count);
    if (size_t(count) != bytes) LOG_FATAL("MpiComm::try_recv: expected " << bytes << " B from rank " << src << " tag " << tag << ", got " << count);
    MPI_Recv(buf, count, MPI_BYTE, src, fold(tag), comm_, MPI_STATUS_IGNORE);
    return true;
  }

  void barrier() override { MPI_Barrier(comm_); }
  void allgather(const void *sendbuf, void *recvbuf, size_t bytesPerRank) override {
    MPI_Allgather(sendbuf, int(bytesPerRank), MPI_BYTE, recvbuf, int(bytesPerRank), MPI_BYTE, comm_);
  }
  void bcast(void *buf, size_t bytes, int root) override { MPI_Bcast(buf, int(bytes), MPI_BYTE, root, comm_); }
  void allgatherv(const void *sendbuf, void *recvbuf, const std::vector<size_t> &counts) override {
    std::vector<int> c(counts.size()), d(counts.size());
    int off = 0;
    for (size_t r = 0; r < counts.size(); ++r) {
      c[r] = int(counts[r]);
      d[r] = off;
      off += c[r];
    }
    MPI_Allgatherv(sendbuf, c[size_t(rank_)], MPI_BYTE, recvbuf, c.data(), d.data(), MPI_BYTE, comm_);
  }

private:
  struct Pending {
    std::vector<char> data;
    MPI_Request req;
    bool inPlace = false; // the caller's own (device) buffer is on the wire
  };
  /* library tags can exceed the smallest legal MPI_TAG_UB; fold the high bits down (collisions would need two live
     messages between the same pair whose tags differ only above bit 14 and alias after folding: the kinds in
     tx_common.hpp occupy the low bits, so distinct kinds never alias) */
  int fold(int tag) const { return tag <= tagUb_ ? tag : int((unsigned(tag) ^ (unsigned(tag) >> 15)) & 0x7fffu); }
  void reap() {
    for (auto it = pending_.begin(); it != pending_.end();) {
      int done = 0;
      MPI_Test(&it->req, &done, MPI_STATUS_IGNORE);
      it = done ? pending_.erase(it) : ++it;
    }
  }

  MPI_Comm comm_;
  int rank_ = 0, size_ = 1, tagUb_ = 32767;
  bool owner_ = false;
  std::list<Pending> pending_;
};

} // namespace

std::unique_ptr<Comm> make_mpi_comm() { return std::unique_ptr<Comm>(new MpiComm()); }

} // namespace stencil

#else

namespace stencil {
std::unique_ptr<Comm> make_mpi_comm() { return nullptr; }
} // namespace stencil

#endif
