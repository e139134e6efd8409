#pragma once
/* Individually constructible transports: one sender / recver pair per (source subdomain, destination subdomain),
   usable without a DistributedDomain — what the reference's transport tests build by hand
   (test/test_cuda_mpi_colocatedtx.cu:43-44 rings of sender/recver objects).

   Parity: reference include/stencil/tx_cuda.cuh
     :513-772  RemoteSender / RemoteRecver : StatefulSender / StatefulRecver — pack, device -> pinned host, MPI_Isend;
               MPI_Irecv, pinned host -> device, unpack; driven by  send()/recv();  while (active()) if (next_ready()) next();  wait();
     :776-974  CudaAwareMpiSender / Recver — the same without the host hop when the communicator takes device pointers.
   Here both are one class each: the hop is dropped when Comm::device_buffers() says the process group takes device
   pointers (MpiComm with -DUSE_CUDA_AWARE_MPI=ON, or a user-supplied communicator).

   DistributedDomain does not use these objects — its Exchanger plans all messages of a subdomain into a few launches —
   they exist for users who wire transports themselves and for tests of the message format (packed layout identical to
   DevicePacker's, tag = make_tag(dstDom, srcDom-in-payload...) as below). */
#include <vector>

#include "stencil/comm.hpp"
#include "stencil/packer.cuh"
#include "stencil/rcstream.hpp"
#include "stencil/tx_common.hpp"

namespace stencil {
namespace detail {
/* one tag per ordered (source subdomain, destination subdomain) pair of two ranks */
inline int remote_tag(int srcDom, int dstDom) { return (3 << 24) | ((srcDom & 0xFF) << 8) | (dstDom & 0xFF); }
} // namespace detail
} // namespace stencil

class RemoteSender : public StatefulSender {
  enum class State { Idle, Staging, Done };
  int srcRank_, srcDom_, dstRank_, dstDom_;
  LocalDomain *domain_;
  stencil::Comm &comm_;
  RcStream stream_;
  DevicePacker packer_;
  char *host_ = nullptr;
  cudaEvent_t staged_ = nullptr;
  State state_ = State::Idle;

public:
  RemoteSender(int srcRank, int srcDom, int dstRank, int dstDom, LocalDomain &domain, stencil::Comm &comm = stencil::world())
      : srcRank_(srcRank), srcDom_(srcDom), dstRank_(dstRank), dstDom_(dstDom), domain_(&domain), comm_(comm), stream_(domain.gpu()),
        packer_(stream_) {}
  ~RemoteSender() override {
    if (staged_) cudaEventDestroy(staged_);
    if (host_) cudaFreeHost(host_);
  }
  RemoteSender(const RemoteSender &) = delete;
  RemoteSender &operator=(const RemoteSender &) = delete;

  void start_prepare(LocalDomain *domain, const std::vector<Message> &outbox) override {
    domain_ = domain;
    packer_.prepare(domain, outbox);
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(domain->gpu()));
    if (!staged_) CUDA_RUNTIME(cudaEventCreateWithFlags(&staged_, cudaEventDisableTiming));
    if (!comm_.device_buffers()) {
      if (host_) CUDA_RUNTIME(cudaFreeHost(host_));
      CUDA_RUNTIME(cudaHostAlloc(&host_, size_t(packer_.size()), cudaHostAllocDefault));
    }
    CUDA_RUNTIME(cudaSetDevice(prev));
  }
  void finish_prepare() override {}

  /* enqueue pack (+ device -> host copy) on this sender's stream; the message leaves in next() */
  void send() override {
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(domain_->gpu()));
    packer_.pack();
    if (host_) CUDA_RUNTIME(cudaMemcpyAsync(host_, packer_.data(), size_t(packer_.size()), cudaMemcpyDeviceToHost, stream_));
    CUDA_RUNTIME(cudaEventRecord(staged_, stream_));
    CUDA_RUNTIME(cudaSetDevice(prev));
    state_ = State::Staging;
  }
  bool active() override { return state_ == State::Staging; }
  bool next_ready() override {
    if (state_ != State::Staging) return false;
    const cudaError_t q = cudaEventQuery(staged_);
    if (q == cudaErrorNotReady) return false;
    CUDA_RUNTIME(q);
    return true;
  }
  void next() override {
    comm_.send(dstRank_, stencil::detail::remote_tag(srcDom_, dstDom_), host_ ? static_cast<const void *>(host_) : packer_.data(),
               size_t(packer_.size()));
    state_ = State::Done;
  }
  /* the packed image may be overwritten by the next send() once this returns */
  void wait() override {
    comm_.flush_sends();
    state_ = State::Idle;
  }
  int64_t size() { return packer_.size(); }
  int src_rank() const { return srcRank_; }
};

class RemoteRecver : public StatefulRecver {
  enum class State { Idle, WaitMessage, Unpacking };
  int srcRank_, srcDom_, dstRank_, dstDom_;
  LocalDomain *domain_;
  stencil::Comm &comm_;
  RcStream stream_;
  DeviceUnpacker unpacker_;
  char *host_ = nullptr;
  State state_ = State::Idle;

public:
  RemoteRecver(int srcRank, int srcDom, int dstRank, int dstDom, LocalDomain &domain, stencil::Comm &comm = stencil::world())
      : srcRank_(srcRank), srcDom_(srcDom), dstRank_(dstRank), dstDom_(dstDom), domain_(&domain), comm_(comm), stream_(domain.gpu()),
        unpacker_(stream_) {}
  ~RemoteRecver() override {
    if (host_) cudaFreeHost(host_);
  }
  RemoteRecver(const RemoteRecver &) = delete;
  RemoteRecver &operator=(const RemoteRecver &) = delete;

  void start_prepare(LocalDomain *domain, const std::vector<Message> &inbox) override {
    domain_ = domain;
    unpacker_.prepare(domain, inbox);
    if (!comm_.device_buffers()) {
      if (host_) CUDA_RUNTIME(cudaFreeHost(host_));
      CUDA_RUNTIME(cudaHostAlloc(&host_, size_t(unpacker_.size()), cudaHostAllocDefault));
    }
  }
  void finish_prepare() override {}

  void recv() override { state_ = State::WaitMessage; }
  bool active() override { return state_ == State::WaitMessage; }
  /* true once the message has arrived (it is then already in the pinned buffer / the unpacker's device buffer) */
  bool next_ready() override {
    if (state_ != State::WaitMessage) return false;
    return comm_.try_recv(srcRank_, stencil::detail::remote_tag(srcDom_, dstDom_), host_ ? static_cast<void *>(host_) : unpacker_.data(),
                          size_t(unpacker_.size()));
  }
  void next() override {
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(domain_->gpu()));
    if (host_) CUDA_RUNTIME(cudaMemcpyAsync(unpacker_.data(), host_, size_t(unpacker_.size()), cudaMemcpyHostToDevice, stream_));
    unpacker_.unpack();
    CUDA_RUNTIME(cudaSetDevice(prev));
    state_ = State::Unpacking;
  }
  /* the halos are in place once this returns */
  void wait() override {
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(domain_->gpu()));
    CUDA_RUNTIME(cudaStreamSynchronize(stream_));
    CUDA_RUNTIME(cudaSetDevice(prev));
    state_ = State::Idle;
  }
  int64_t size() { return unpacker_.size(); }
  int dst_rank() const { return dstRank_; }
};
