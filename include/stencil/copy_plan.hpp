#pragma once
/* CopyPlan: the one data-movement primitive of this library.

   Every pack / unpack / translate / halo-push in the reference (kernels K1-K7, SURVEY 2.3: pack_kernel.cuh,
   copy.cuh, packer.cuh:52-69,194-250) is "copy a 3-D box between two pitched arrays". Here all of them are
   *segments* of one table executed by ONE launch of box_copy_kernel (src/box_copy.cu): a segment is a box of one
   quantity with byte pitches on both sides, and a chunk table splits segments into ~equal pieces of work so a thin
   corner and a 3 MB face keep the same number of bytes per CTA. A dense pack buffer is simply a view whose pitches
   equal the box extent; a neighbour GPU's halo is a view whose base pointer is a peer address.

   Both endpoints carry two base pointers indexed by parity so that LocalDomain::swap() (curr <-> next) needs no
   device-side update and no re-planning: the launch passes the current parity. */
#include <cstddef>
#include <cstdint>
#include <vector>

#include <cuda_runtime.h>

#include "stencil/dim3.hpp"

namespace stencil {

struct CopySeg {
  const char *src[2]; // first byte of the box, per parity
  char *dst[2];
  long long srcPitchY, srcPitchZ; // bytes between consecutive y rows / z planes
  long long dstPitchY, dstPitchZ;
  unsigned rowBytes; // contiguous bytes per row (extent.x * elemSize)
  unsigned ny, nz;   // rows per plane, planes
  unsigned vecLog2;  // log2 of the widest access (bytes) legal for every row of both sides: 0..4
  /* "sector mode" for thin rows (x faces, edges, corners) whose destination row lies next to row padding: the row
     is written as whole 32-byte sectors — padLoWords words of padding, the payload, padding up to totalWords — so the
     receiving memory system never has to read-modify-write a partially written sector. 0 = plain mode. */
  unsigned short padLoWords, totalWords;
  /* 1: thin rows (<= 64 B) going to a dense destination (rows back to back): gathered through shared memory and
     written as one contiguous, fully coalesced run — what a remote mailbox wants */
  unsigned gatherDense;
  /* thin rows of at most 16 bytes whose enclosing aligned 16-byte windows may be read (the source is a padded
     LocalDomain buffer) and start at the same word offset in every row: 1 + that offset (0..3), else 0. The payload is
     then fetched with one or two 16-byte loads instead of a 4-byte load per word — thin rows are bound by the number of
     memory requests, not by bytes (tools/microbench/xface.cu: 91 -> 66 us for 1 M rows). */
  unsigned srcWin;
};

struct CopyChunk {
  unsigned seg;   // index into the segment table
  unsigned z;     // first plane of the box this chunk covers
  unsigned y0;    // first row inside that plane (0 when the chunk spans several planes)
  unsigned short nrows;   // rows per plane in this chunk
  unsigned short nplanes; // whole planes covered (1: a run of rows inside one plane — no index division on device)
};

/* a strided 3-D array: address of element (0,0,0) per parity, pitches in bytes */
struct BoxView {
  char *base[2];
  long long pitchY, pitchZ;
  size_t elemSize;
  /* bytes of don't-care padding before logical x = 0 and after logical x = rowElems of EVERY row (0: none / unknown) */
  long long padLoBytes = 0, padHiBytes = 0, rowElems = 0;
  /* destination lives behind NVLink: few large write packets matter more than instruction count (thin rows are then
     gathered through shared memory and written as contiguous runs) */
  bool acrossLink = false;
  /* every aligned 16-byte window that overlaps a row of this array lies inside its allocation (rows start on 16-byte
     boundaries, pitches are multiples of 16 bytes, the buffer has slack after its last row) */
  bool windowSafe = false;
  static BoxView dense(void *p, const Dim3 &ext, size_t elemSize) {
    BoxView v;
    v.base[0] = v.base[1] = static_cast<char *>(p);
    v.pitchY = (long long)(ext.x) * (long long)elemSize;
    v.pitchZ = v.pitchY * ext.y;
    v.elemSize = elemSize;
    return v;
  }
  static BoxView pitched(void *p0, void *p1, const Dim3 &pitchElems, size_t elemSize) {
    BoxView v;
    v.base[0] = static_cast<char *>(p0);
    v.base[1] = static_cast<char *>(p1);
    v.pitchY = (long long)(pitchElems.x) * (long long)elemSize;
    v.pitchZ = v.pitchY * pitchElems.y;
    v.elemSize = elemSize;
    return v;
  }
};

/* remote (or local) 32-bit epoch flags written after a system-scope fence once every byte of the plan has been stored */
struct SignalSet {
  static constexpr int kMax = 32;
  unsigned n;
  unsigned *flag[kMax];
};

class CopyPlan {
  std::vector<CopySeg> segs_;
  std::vector<CopyChunk> chunks_;
  size_t bytes_ = 0;
  CopySeg *dSegs_ = nullptr;
  CopyChunk *dChunks_ = nullptr;
  unsigned *dDone_ = nullptr; // CTA completion counter used by the signalling variant
  int dev_ = -1;
  std::vector<unsigned char> optional_; // per segment
  std::vector<unsigned char> remote_;   // per segment: the destination is on another GPU
  unsigned nRequiredChunks_ = 0;

  void upload();

public:
  CopyPlan() {}
  ~CopyPlan();
  CopyPlan(const CopyPlan &) = delete;
  CopyPlan &operator=(const CopyPlan &) = delete;
  CopyPlan(CopyPlan &&o) noexcept { swap(o); }
  CopyPlan &operator=(CopyPlan &&o) noexcept {
    swap(o);
    return *this;
  }
  void swap(CopyPlan &o) noexcept;

  /* copy box `ext` (elements) from src at srcPos to dst at dstPos; zero-volume boxes are ignored */
  void add(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext);
  /* same, but launch_and_signal(..., withOptional = false) leaves these boxes out */
  void add_optional(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext);
  void clear();

  size_t num_segments() const { return segs_.size(); }
  size_t num_chunks() const { return chunks_.size(); }
  size_t bytes() const { return bytes_; } // payload bytes moved per launch
  bool empty() const { return segs_.empty(); }
  const std::vector<CopySeg> &segments() const { return segs_; }

  /* build the chunk table and upload both tables to device `dev` (the device that will run the kernel) */
  void finalize(int dev, size_t chunkBytes = 32 * 1024);
  bool finalized() const { return dSegs_ != nullptr || segs_.empty(); }
  /* the host half of finalize(): the chunk table alone (needs no GPU; exposed for tests and tools) */
  void build_chunks(size_t chunkBytes = 32 * 1024);
  const std::vector<CopyChunk> &chunks() const { return chunks_; }
  size_t num_required_chunks() const { return nRequiredChunks_; }

  /* one launch on `stream` (a stream of the plan's device); no host synchronisation */
  void launch(int parity, cudaStream_t stream) const;
  /* as launch(), then the last CTA to finish publishes `epoch` to every flag of `sig` (release, system scope) */
  void launch_and_signal(int parity, const SignalSet &sig, unsigned epoch, cudaStream_t stream, bool withOptional = true) const;
  /* number of CTAs a launch uses */
  unsigned grid_size(bool withOptional = true) const;
};

/* Spin (one warp, bounded by timeoutNs) until every flag has reached `epoch` (acquire, system scope). On timeout
   writes 1+index of the late flag to *errWord (host-mapped) and returns, so a lost peer cannot hang the GPU.
   Never use this on flags that a kernel on the SAME device signals — use stream events there. */
struct WaitSet {
  static constexpr int kMax = 32;
  unsigned n;
  const unsigned *flag[kMax];
};
void launch_wait_flags(const WaitSet &ws, unsigned epoch, unsigned *errWord, unsigned long long timeoutNs,
                       cudaStream_t stream);
/* how long a device-side flag wait may last before it reports through the error word: STENCIL_FLAG_TIMEOUT_MS
   (default 5000). One helper for the Exchanger and both engines, so that debuggers / MPS / slow peers need one knob. */
unsigned long long flag_timeout_ns();
/* publish `epoch` to every flag (release, system scope) once all prior work of `stream` is done */
void launch_signal_flags(const SignalSet &sig, unsigned epoch, cudaStream_t stream);

} // namespace stencil

namespace stencil {
/* Table-free launches for a handful of boxes (descriptors travel as kernel parameters): used by the standalone
   pack/unpack/translate entry points where building a CopyPlan would cost more than the copy. */
struct SegBatch {
  static constexpr int kMax = 8;
  unsigned n;
  CopySeg seg[kMax];
};
/* make a segment descriptor without a plan */
CopySeg make_seg(const BoxView &src, const Dim3 &srcPos, const BoxView &dst, const Dim3 &dstPos, const Dim3 &ext);
void launch_box_copy(const SegBatch &batch, int parity, cudaStream_t stream);
} // namespace stencil
