#pragma once
/* LocalDomain: one subdomain on one GPU — N untyped quantities, each with a `curr` and a `next` buffer that
   include the halo, plus the halo geometry queries the exchange is planned from.

   Public surface follows the reference (include/stencil/local_domain.cuh:17-349, src/local_domain.cu:1-169):
   DataHandle<T>, add_data, set_radius, get_curr/get_next, curr_data/next_data, get_{curr,next}_accessor,
   get_compute_region/get_full_region, halo_pos/halo_coords/halo_extent/halo_bytes, size/raw_size/origin/gpu, swap,
   region_to_host/interior_to_host/quantity_to_host, realize, dev_curr_datas/dev_elem_sizes.

   B200-first storage (differs from the reference's one dense cudaMalloc per buffer):
   * ONE slab per subdomain holds every quantity's curr+next and the arrival flags, so a single CUDA IPC handle (or
     peer mapping) exposes the whole subdomain to the 7 other GPUs of the NVSwitch box;
   * rows are padded: the first interior element (x = 0) of every row sits on a `alignment()` (default 128 B)
     boundary and the row pitch is a multiple of it — legal for TMA tensor maps (16 B strides) and makes y/z-face
     traffic 16 B-vectorisable. raw_size() stays the LOGICAL extent (sz + radii); pitch(qi) is the real one, and
     Accessor carries it, so kernels written against Accessor do not notice;
   * swap() is pointer parity only: no cudaMemcpy, no synchronisation (the reference blocks on a memcpy per call).
   LocalDomain is move-only (it owns device memory; the reference copies raw owning pointers around). */
#include <cstdint>
#include <string>
#include <vector>

#include "stencil/accessor.hpp"
#include "stencil/copy_plan.hpp"
#include "stencil/cuda_runtime.hpp"
#include "stencil/dim3.hpp"
#include "stencil/halo_geometry.hpp"
#include "stencil/radius.hpp"
#include "stencil/rect3.hpp"

class DistributedDomain;

template <typename T> class DataHandle {
  friend class DistributedDomain;
  friend class LocalDomain;
  size_t id_;
  std::string name_;

public:
  DataHandle(size_t i = 0, const std::string &name = "") : id_(i), name_(name) {}
  size_t id() const noexcept { return id_; }
  const std::string &name() const noexcept { return name_; }
};

class LocalDomain {
  friend class DistributedDomain;

public:
  /* words of the per-subdomain control block at the start of the slab */
  static constexpr size_t kArriveFlags = 27; // one arrival epoch per source direction slot
  static constexpr size_t kControlBytes = 512;

private:
  Dim3 sz_;
  Dim3 origin_;
  Radius radius_;
  int dev_;
  size_t align_ = 128;

  std::vector<size_t> elemSize_;
  std::vector<std::string> name_;

  struct Layout {
    int64_t pitchX = 0; // elements per row (padded)
    int64_t lead = 0;   // padding elements before logical (0,0,0) so that interior x=0 is aligned
    size_t bufBytes = 0;
    size_t off[2] = {0, 0}; // slab offsets of the two buffers
  };
  std::vector<Layout> layout_;

  /* x-side mailboxes (two parities): dense landing areas for halo messages whose rows are only a few cells long.
     A remote GPU writes such a message here as one contiguous block (NVLink moves ~5 G packets/s whatever their
     size, so 12-byte rows sent one by one reach a fifth of the link rate) and the owner scatters it into its halo. */
  std::vector<size_t> mboxOff_; // [slot(side) * nq + qi], relative to the mailbox base; sides with x == 0 unused
  size_t mboxBytes_ = 0;        // one parity
  size_t mboxBase_[2] = {0, 0}; // slab offsets

  char *slab_ = nullptr;
  size_t slabBytes_ = 0;
  bool ownsSlab_ = false;
  int parity_ = 0; // curr = buffer[parity_], next = buffer[parity_ ^ 1]

  void **devPtrTable_[2] = {nullptr, nullptr}; // device arrays of curr pointers, one per parity
  size_t *devElemSizes_ = nullptr;

  void release();

public:
  LocalDomain(Dim3 sz, Dim3 origin, int dev) : sz_(sz), origin_(origin), dev_(dev) {}
  ~LocalDomain() { release(); }
  LocalDomain(const LocalDomain &) = delete;
  LocalDomain &operator=(const LocalDomain &) = delete;
  LocalDomain(LocalDomain &&o) noexcept;
  LocalDomain &operator=(LocalDomain &&o) noexcept;

  void set_device(CudaErrorsFatal fatal = CudaErrorsFatal::YES);

  int64_t num_data() const { return int64_t(elemSize_.size()); }
  const Dim3 &origin() const noexcept { return origin_; }

  /* add an untyped quantity of n bytes per element; returns its index. Only before realize(). */
  int64_t add_data(size_t n, const std::string &name = "");
  template <typename T> DataHandle<T> add_data(const std::string &name = "") {
    return DataHandle<T>(size_t(add_data(sizeof(T), name)), name);
  }

  void set_radius(size_t r) { radius_ = Radius::constant(r); }
  void set_radius(const Radius &r) { radius_ = r; }
  const Radius &radius() const noexcept { return radius_; }

  /* row alignment in bytes (power of two); only before realize() */
  void set_alignment(size_t bytes) { align_ = bytes; }
  size_t alignment() const noexcept { return align_; }

  size_t elem_size(size_t qi) const { return elemSize_.at(qi); }
  const std::string &data_name(size_t qi) const { return name_.at(qi); }

  /* pointer to logical element (0,0,0) — the corner of the negative-side halo */
  void *curr_data(size_t qi) const { return buffer(qi, parity_); }
  void *next_data(size_t qi) const { return buffer(qi, parity_ ^ 1); }
  /* buffer `which` (0/1) regardless of the current parity */
  void *buffer(size_t qi, int which) const {
    const Layout &l = layout_.at(qi);
    return slab_ ? slab_ + l.off[which] + size_t(l.lead) * elemSize_[qi] : nullptr;
  }
  template <typename T> T *get_curr(const DataHandle<T> &h) const { return static_cast<T *>(curr_data(h.id_)); }
  template <typename T> T *get_next(const DataHandle<T> &h) const { return static_cast<T *>(next_data(h.id_)); }

  /* device-resident table of curr pointers / element sizes (API parity; the table tracks swap()) */
  void **dev_curr_datas() const { return devPtrTable_[parity_]; }
  size_t *dev_elem_sizes() const { return devElemSizes_; }

  /* real pitch of quantity qi in elements: (padded row, rows per plane, planes) */
  Dim3 pitch(size_t qi) const { return Dim3(layout_.at(qi).pitchX, raw_size().y, raw_size().z); }

  template <typename T> Accessor<T> get_curr_accessor(const DataHandle<T> &h) const noexcept {
    return Accessor<T>(get_curr(h), halo_origin(), pitch(h.id_));
  }
  template <typename T> Accessor<T> get_next_accessor(const DataHandle<T> &h) const noexcept {
    return Accessor<T>(get_next(h), halo_origin(), pitch(h.id_));
  }

  /* global coordinate of logical element (0,0,0) */
  Dim3 halo_origin() const noexcept {
    return Dim3(origin_.x - int64_t(radius_.x(-1)), origin_.y - int64_t(radius_.y(-1)), origin_.z - int64_t(radius_.z(-1)));
  }

  Rect3 get_compute_region() const noexcept { return Rect3(origin_, origin_ + sz_); }
  Rect3 get_full_region() const noexcept { return Rect3(halo_origin(), halo_origin() + raw_size()); }

  /* offset (elements, from logical (0,0,0)) of the region on side `dir`: the halo itself, or the adjacent compute
     cells when halo == false. dir = 0 gives the compute region. */
  Dim3 halo_pos(const Dim3 &dir, bool halo) const noexcept { return stencil::halo_pos(dir, sz_, radius_, halo); }
  Rect3 halo_coords(const Dim3 &dir, bool halo) const {
    const Dim3 lo = halo_pos(dir, halo) + halo_origin();
    return Rect3(lo, lo + halo_extent(dir));
  }
  static Dim3 halo_extent(const Dim3 &dir, const Dim3 &sz, const Radius &radius) {
    return stencil::halo_extent(dir, sz, radius);
  }
  Dim3 halo_extent(const Dim3 &dir) const noexcept { return stencil::halo_extent(dir, sz_, radius_); }
  int64_t halo_bytes(const Dim3 &dir, int64_t qi) const noexcept {
    return int64_t(elemSize_[size_t(qi)]) * halo_extent(dir).flatten();
  }

  Dim3 size() const noexcept { return sz_; }
  /* logical allocation extent in elements (compute region + halos); see pitch() for the padded row length */
  Dim3 raw_size() const noexcept {
    return Dim3(sz_.x + int64_t(radius_.x(-1) + radius_.x(1)), sz_.y + int64_t(radius_.y(-1) + radius_.y(1)),
                sz_.z + int64_t(radius_.z(-1) + radius_.z(1)));
  }
  int gpu() const { return dev_; }
  int parity() const noexcept { return parity_; }

  /* exchange curr and next: O(1), host only */
  void swap() noexcept { parity_ ^= 1; }

  /* strided view of quantity qi (both parities: base[p] is the curr buffer when parity() == p) for CopyPlan */
  stencil::BoxView view(size_t qi) const { return view_at(qi, slab_); }
  /* the same view for a copy of this subdomain's slab mapped at `slabBase` (e.g. through CUDA IPC in a peer process) */
  stencil::BoxView view_at(size_t qi, char *slabBase) const {
    const Layout &l = layout_.at(qi);
    const long long es = (long long)elemSize_[qi];
    char *b0 = slabBase ? slabBase + l.off[0] + size_t(l.lead) * elemSize_[qi] : nullptr;
    char *b1 = slabBase ? slabBase + l.off[1] + size_t(l.lead) * elemSize_[qi] : nullptr;
    stencil::BoxView v = stencil::BoxView::pitched(b0, b1, pitch(qi), elemSize_[qi]);
    /* the cells between the end of one logical row and the start of the next are padding (the buffer ends with 32
       bytes of slack for the last row): movers may overwrite up to one sector's worth of it */
    v.rowElems = raw_size().x;
    v.padLoBytes = l.lead * es < 28 ? l.lead * es : 28;
    v.padHiBytes = (l.pitchX - raw_size().x) * es < 28 ? (l.pitchX - raw_size().x) * es : 28;
    /* aligned layouts only (compute_layout's first branch): buffer starts and row pitches are multiples of 128 bytes */
    v.windowSafe = (v.pitchY % 16) == 0 && (v.pitchZ % 16) == 0 && ((l.off[0] | l.off[1]) % 16) == 0;
    return v;
  }

  /* dense copy of box [pos, pos+ext) (elements from logical (0,0,0)) of curr quantity qi */
  std::vector<unsigned char> region_to_host(const Dim3 &pos, const Dim3 &ext, size_t qi) const;
  std::vector<unsigned char> interior_to_host(size_t qi) const {
    return region_to_host(halo_pos(Dim3(0, 0, 0), true), halo_extent(Dim3(0, 0, 0)), qi);
  }
  std::vector<unsigned char> quantity_to_host(size_t qi) const { return region_to_host(Dim3(0, 0, 0), raw_size(), qi); }
  /* inverse of region_to_host: scatter a dense host box into curr quantity qi */
  void region_from_host(const void *host, const Dim3 &pos, const Dim3 &ext, size_t qi);

  /* allocate. Idempotent per object; add_data/set_radius must come first. */
  void realize();
  bool realized() const noexcept { return slab_ != nullptr; }

  /* slab access for peer/IPC mapping */
  void *slab() const noexcept { return slab_; }
  size_t slab_bytes() const noexcept { return slabBytes_; }
  size_t buffer_offset(size_t qi, int which) const {
    return layout_.at(qi).off[which] + size_t(layout_[qi].lead) * elemSize_[qi];
  }
  /* Mailbox of the halo on side `side` (side.x != 0) of quantity qi: slab offset of a dense halo_extent(side) box.
     `parity` 0/1 alternates per exchange / iteration so a sender can fill one while the owner drains the other. */
  size_t mailbox_offset(const Dim3 &side, size_t qi, int parity) const {
    return mboxBase_[parity] + mboxOff_.at(size_t(DirectionMap<int>::slot(side)) * elemSize_.size() + qi);
  }
  size_t mailbox_bytes() const noexcept { return mboxBytes_; }
  /* first byte of the padded buffer (row 0 starts here, `lead_elems` before logical element (0,0,0)) */
  size_t buffer_start_offset(size_t qi, int which) const { return layout_.at(qi).off[which]; }
  int64_t lead_elems(size_t qi) const { return layout_.at(qi).lead; }
  /* arrival-flag array (kArriveFlags x uint32) at the start of the slab */
  unsigned *arrive_flags() const noexcept { return reinterpret_cast<unsigned *>(slab_); }

  /* layout computation only (no allocation): used by peers to derive offsets inside a mapped remote slab */
  void compute_layout();
};
