#pragma once
/* Message descriptors, message tags and the stateful transport interfaces.
   Parity: reference include/stencil/tx_common.hpp:8-177 — Message{dir_, srcGPU_, dstGPU_} ordered by direction
   only; MsgKind; make_tag<kind>(payload, dir) with bits 0-1 kind, 2-7 direction, 8-23 payload (24-bit budget kept
   so tags stay valid for MPIs whose tag ceiling is 2^23); legacy make_tag(gpu, idx, dir); StatefulSender/Recver. */
#include <cassert>
#include <vector>

#include "stencil/dim3.hpp"

class Message {
public:
  Dim3 dir_;
  int srcGPU_; // subdomain id within the sending rank
  int dstGPU_; // subdomain id within the receiving rank
  Message(const Dim3 &dir, int srcGPU, int dstGPU) : dir_(dir), srcGPU_(srcGPU), dstGPU_(dstGPU) {}
  bool operator<(const Message &r) const noexcept { return dir_ < r.dir_; }
  bool operator==(const Message &r) const noexcept { return dir_ == r.dir_ && srcGPU_ == r.srcGPU_ && dstGPU_ == r.dstGPU_; }
};

enum class MsgKind { ColocatedEvt = 0, ColocatedMem = 1, ColocatedDev = 2, ColocatedNotify = 3, Other = 4 };

namespace stencil {
namespace detail {
/* 2 bits per axis: 0 -> 00, +1 -> 01, -1 -> 10; x in the low bits */
inline int dir_bits(const Dim3 &d) {
  auto enc = [](int64_t v) { return v == 0 ? 0 : (v == 1 ? 1 : 2); };
  assert(d.all_gt(-2) && d.all_lt(2));
  return enc(d.x) | (enc(d.y) << 2) | (enc(d.z) << 4);
}
} // namespace detail
} // namespace stencil

template <MsgKind kind> inline int make_tag(int payload, const Dim3 &dir = Dim3(0, 0, 0)) {
  static_assert(static_cast<int>(kind) >= 0 && static_cast<int>(kind) <= 3, "kind must fit in two bits");
  assert(payload >= 0 && payload < (1 << 16));
  return static_cast<int>(kind) | (stencil::detail::dir_bits(dir) << 2) | ((payload & 0xFFFF) << 8);
}

/* legacy tag: data index bits 0-15, gpu id bits 16-23, direction bits 24-29 */
inline int make_tag(int gpu, int idx, const Dim3 &dir) {
  assert(gpu >= 0 && gpu < (1 << 8) && idx >= 0 && idx < (1 << 16));
  return (idx & 0xFFFF) | ((gpu & 0xFF) << 16) | (stencil::detail::dir_bits(dir) << 24);
}

/* A transport with phases:  send();  while (active()) if (next_ready()) next();  wait(); */
class StatefulSender {
public:
  virtual ~StatefulSender() {}
  /* allocate buffers / build plans for these outgoing messages of `domain` */
  virtual void start_prepare(class LocalDomain *domain, const std::vector<Message> &outbox) = 0;
  virtual void finish_prepare() = 0;
  virtual void send() = 0;
  virtual bool active() = 0;
  virtual bool next_ready() = 0;
  virtual void next() = 0;
  virtual void wait() = 0;
};

class StatefulRecver {
public:
  virtual ~StatefulRecver() {}
  virtual void start_prepare(class LocalDomain *domain, const std::vector<Message> &inbox) = 0;
  virtual void finish_prepare() = 0;
  virtual void recv() = 0;
  virtual bool active() = 0;
  virtual bool next_ready() = 0;
  virtual void next() = 0;
  virtual void wait() = 0;
};
