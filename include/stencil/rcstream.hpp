#pragma once
/* RcStream: reference-counted RAII cudaStream_t bound to a device.
   Parity: reference include/stencil/rcstream.hpp:9-72, src/rcstream.cpp:5-46 — Priority::{DEFAULT,HIGH},
   non-blocking stream, copy/move share one stream that is destroyed with the last owner, implicit conversion to
   cudaStream_t, device(). Implementation here is a shared_ptr with a deleter rather than a hand-rolled counter. */
#include <memory>

#include <cuda_runtime.h>

class RcStream {
public:
  enum class Priority { DEFAULT, HIGH };

  explicit RcStream(int dev, Priority p = Priority::DEFAULT);
  /* new DEFAULT-priority stream on device 0 (reference behaviour) */
  RcStream() : RcStream(0) {}

  operator cudaStream_t() const noexcept { return h_ ? h_->stream : nullptr; }
  cudaStream_t get() const noexcept { return h_ ? h_->stream : nullptr; }
  int device() const noexcept { return h_ ? h_->dev : -1; }
  /* number of RcStream objects sharing this stream */
  long use_count() const noexcept { return h_.use_count(); }
  bool operator==(const RcStream &r) const noexcept { return get() == r.get(); }
  bool operator!=(const RcStream &r) const noexcept { return !(*this == r); }
  void synchronize() const;

private:
  struct Handle {
    cudaStream_t stream;
    int dev;
  };
  std::shared_ptr<Handle> h_;
};
