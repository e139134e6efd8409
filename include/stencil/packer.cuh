#pragma once
/* DevicePacker / DeviceUnpacker: gather all outgoing (scatter all incoming) halo messages of one subdomain for one
   peer into (out of) ONE dense device buffer.

   Parity: reference include/stencil/packer.cuh:22-48 (Packer/Unpacker), :71-192 (DevicePacker), :252-364
   (DeviceUnpacker). Buffer layout is identical — messages sorted by direction (x-major Dim3 order); inside a message
   the quantities in index order, each aligned to its element size; a message sent towards +d is sized by the
   RECEIVER's -d halo (halo_bytes(dir * -1)) — so packed buffers are wire-compatible with the reference layout
   (e.g. radius +x=2/-x=1 with float,char,double on 10x10x10... gives the same 264 B as test_cuda_packer.cu:74-92).

   B200-first execution: the whole buffer is ONE CopyPlan launch (all messages x all quantities, 16 B vectors where
   the rows allow) instead of one captured kernel per message that loops serially over quantities; curr/next parity
   is a launch argument so plans survive swap() with no device pointer table. Buffers are freed (the reference
   leaks them). */
#include <algorithm>
#include <vector>

#include "stencil/align.cuh"
#include "stencil/copy_plan.hpp"
#include "stencil/local_domain.cuh"
#include "stencil/logging.hpp"
#include "stencil/tx_common.hpp"

class Packer {
public:
  virtual ~Packer() {}
  virtual void prepare(LocalDomain *domain, const std::vector<Message> &messages) = 0;
  virtual void pack() = 0;
  virtual int64_t size() = 0;
  virtual void *data() = 0;
};

class Unpacker {
public:
  virtual ~Unpacker() {}
  virtual void prepare(LocalDomain *domain, const std::vector<Message> &messages) = 0;
  virtual void unpack() = 0;
  virtual int64_t size() = 0;
  virtual void *data() = 0;
};

namespace stencil {
namespace detail {

/* byte offset of every (message, quantity) block in the packed buffer, plus the total size */
inline int64_t packed_layout(const LocalDomain &dom, const std::vector<Message> &sorted, std::vector<int64_t> *offsets) {
  int64_t size = 0;
  for (const Message &m : sorted)
    for (int64_t qi = 0; qi < dom.num_data(); ++qi) {
      size = int64_t(next_align_of(size_t(size), dom.elem_size(size_t(qi))));
      if (offsets) offsets->push_back(size);
      size += dom.halo_bytes(m.dir_ * -1, qi);
    }
  return size;
}

/* shared buffer + plan management of packer and unpacker */
class PackBase {
protected:
  LocalDomain *domain_ = nullptr;
  std::vector<Message> msgs_;
  int64_t size_ = -1;
  char *buf_ = nullptr;
  bool ownsBuf_ = false;
  cudaStream_t stream_;
  CopyPlan plan_;

  explicit PackBase(cudaStream_t s) : stream_(s) {}
  ~PackBase() { free_buf(); }
  void free_buf() {
    if (buf_ && ownsBuf_) {
      int prev = 0;
      if (cudaGetDevice(&prev) == cudaSuccess) {
        cudaSetDevice(domain_->gpu());
        cudaFree(buf_);
        cudaSetDevice(prev);
      }
    }
    buf_ = nullptr;
  }
  /* external = place the packed image in caller-owned memory (e.g. a peer-mapped buffer) instead of allocating */
  void setup(LocalDomain *domain, const std::vector<Message> &messages, bool packing, void *external) {
    free_buf();
    domain_ = domain;
    msgs_ = messages;
    std::stable_sort(msgs_.begin(), msgs_.end());
    std::vector<int64_t> offs;
    size_ = packed_layout(*domain, msgs_, &offs);
    if (size_ == 0) LOG_FATAL("zero-size packer was prepared");
    int prev = 0;
    CUDA_RUNTIME(cudaGetDevice(&prev));
    CUDA_RUNTIME(cudaSetDevice(domain->gpu()));
    if (external) {
      buf_ = static_cast<char *>(external);
      ownsBuf_ = false;
    } else {
      CUDA_RUNTIME(cudaMalloc(&buf_, size_t(size_)));
      ownsBuf_ = true;
    }
    plan_.clear();
    size_t k = 0;
    for (const Message &m : msgs_) {
      /* packing: the compute cells on our `dir` side, shaped like the receiver's -dir halo;
         unpacking: a message that travelled in `dir` lands in our halo on side -dir */
      const Dim3 ext = domain->halo_extent(m.dir_ * -1);
      if (ext.flatten() == 0) LOG_FATAL("message " << m.dir_ << " has zero size, ext=" << ext);
      const Dim3 pos = packing ? domain->halo_pos(m.dir_, false) : domain->halo_pos(m.dir_ * -1, true);
      for (int64_t qi = 0; qi < domain->num_data(); ++qi, ++k) {
        const BoxView dense = BoxView::dense(buf_ + offs[k], ext, domain->elem_size(size_t(qi)));
        if (packing)
          plan_.add(domain->view(size_t(qi)), pos, dense, Dim3(0, 0, 0), ext);
        else
          plan_.add(dense, Dim3(0, 0, 0), domain->view(size_t(qi)), pos, ext);
      }
    }
    plan_.finalize(domain->gpu());
    CUDA_RUNTIME(cudaSetDevice(prev));
  }
};

} // namespace detail
} // namespace stencil

class DevicePacker : public Packer, private stencil::detail::PackBase {
public:
  explicit DevicePacker(cudaStream_t stream) : PackBase(stream) {}
  void prepare(LocalDomain *domain, const std::vector<Message> &messages) override { setup(domain, messages, true, nullptr); }
  /* pack straight into caller-provided device memory of at least size() bytes */
  void prepare_into(LocalDomain *domain, const std::vector<Message> &messages, void *buffer) {
    setup(domain, messages, true, buffer);
  }
  /* enqueue on the packer's stream; reads the domain's CURRENT curr buffers */
  void pack() override { plan_.launch(domain_->parity(), stream_); }
  int64_t size() override { return size_; }
  void *data() override { return buf_; }
  const stencil::CopyPlan &plan() const { return plan_; }
  static int64_t packed_size(const LocalDomain &dom, std::vector<Message> messages) {
    std::stable_sort(messages.begin(), messages.end());
    return stencil::detail::packed_layout(dom, messages, nullptr);
  }
};

class DeviceUnpacker : public Unpacker, private stencil::detail::PackBase {
public:
  explicit DeviceUnpacker(cudaStream_t stream) : PackBase(stream) {}
  void prepare(LocalDomain *domain, const std::vector<Message> &messages) override { setup(domain, messages, false, nullptr); }
  void prepare_from(LocalDomain *domain, const std::vector<Message> &messages, void *buffer) {
    setup(domain, messages, false, buffer);
  }
  void unpack() override { plan_.launch(domain_->parity(), stream_); }
  int64_t size() override { return size_; }
  void *data() override { return buf_; }
  const stencil::CopyPlan &plan() const { return plan_; }
};
