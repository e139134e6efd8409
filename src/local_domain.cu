/* LocalDomain storage and host transfers. Reference: src/local_domain.cu:1-169 (see header for the design). */
#include "stencil/nvtx.hpp"
#include "stencil/local_domain.cuh"

#include <cstdlib>
#include <cstring>

#include "stencil/align.cuh"
#include "stencil/logging.hpp"

namespace {
struct DevGuard {
  int prev = 0;
  explicit DevGuard(int d) {
    CUDA_RUNTIME(cudaGetDevice(&prev));
    if (prev != d) CUDA_RUNTIME(cudaSetDevice(d));
  }
  ~DevGuard() { cudaSetDevice(prev); }
};
bool is_pow2(size_t v) { return v && !(v & (v - 1)); }
} // namespace

LocalDomain::LocalDomain(LocalDomain &&o) noexcept : sz_(o.sz_), origin_(o.origin_), radius_(o.radius_), dev_(o.dev_) {
  *this = std::move(o);
}

LocalDomain &LocalDomain::operator=(LocalDomain &&o) noexcept {
  if (this == &o) return *this;
  release();
  sz_ = o.sz_;
  origin_ = o.origin_;
  radius_ = o.radius_;
  dev_ = o.dev_;
  align_ = o.align_;
  elemSize_ = std::move(o.elemSize_);
  name_ = std::move(o.name_);
  layout_ = std::move(o.layout_);
  mboxOff_ = std::move(o.mboxOff_);
  mboxBytes_ = o.mboxBytes_;
  mboxBase_[0] = o.mboxBase_[0];
  mboxBase_[1] = o.mboxBase_[1];
  slab_ = o.slab_;
  slabBytes_ = o.slabBytes_;
  ownsSlab_ = o.ownsSlab_;
  parity_ = o.parity_;
  devPtrTable_[0] = o.devPtrTable_[0];
  devPtrTable_[1] = o.devPtrTable_[1];
  devElemSizes_ = o.devElemSizes_;
  o.slab_ = nullptr;
  o.ownsSlab_ = false;
  o.devPtrTable_[0] = o.devPtrTable_[1] = nullptr;
  o.devElemSizes_ = nullptr;
  return *this;
}

void LocalDomain::release() {
  if (!slab_ && !devPtrTable_[0] && !devElemSizes_) return;
  int prev = 0;
  if (cudaGetDevice(&prev) != cudaSuccess) return; // runtime already torn down
  cudaSetDevice(dev_);
  if (ownsSlab_ && slab_) cudaFree(slab_);
  if (devPtrTable_[0]) cudaFree(devPtrTable_[0]); // both tables live in one allocation
  if (devElemSizes_) cudaFree(devElemSizes_);
  cudaSetDevice(prev);
  slab_ = nullptr;
  devPtrTable_[0] = devPtrTable_[1] = nullptr;
  devElemSizes_ = nullptr;
}

void LocalDomain::set_device(CudaErrorsFatal fatal) {
  const cudaError_t e = cudaSetDevice(dev_);
  if (fatal == CudaErrorsFatal::YES) {
    CUDA_RUNTIME(e);
  } else if (e != cudaSuccess) {
    (void)cudaGetLastError();
  }
}

int64_t LocalDomain::add_data(size_t n, const std::string &name) {
  if (slab_) LOG_FATAL("LocalDomain::add_data after realize()");
  elemSize_.push_back(n);
  name_.push_back(name);
  return int64_t(elemSize_.size()) - 1;
}

void LocalDomain::compute_layout() {
  const Dim3 raw = raw_size();
  layout_.assign(elemSize_.size(), Layout());
  size_t off = kControlBytes;
  const size_t slabAlign = std::max<size_t>(align_, 256);
  for (size_t qi = 0; qi < elemSize_.size(); ++qi) {
    const size_t es = elemSize_[qi];
    Layout &l = layout_[qi];
    const int64_t rmx = int64_t(radius_.x(-1));
    if (is_pow2(es) && is_pow2(align_) && es <= align_) {
      const int64_t per = int64_t(align_ / es); // elements per alignment unit
      const int64_t interiorStart = int64_t(round_up_multiple(size_t(rmx), size_t(per)));
      l.lead = interiorStart - rmx;
      l.pitchX = int64_t(round_up_multiple(size_t(interiorStart + sz_.x + int64_t(radius_.x(1))), size_t(per)));
    } else { // odd element size: dense rows
      l.lead = 0;
      l.pitchX = raw.x;
    }
    /* + one sector of slack: fused kernels write the whole 32-byte sector around a +x halo cell, which for the very
       last row may extend past the last element */
    l.bufBytes = size_t(l.pitchX) * size_t(raw.y) * size_t(raw.z) * es + 32;
    for (int b = 0; b < 2; ++b) {
      off = round_up_multiple(off, slabAlign);
      l.off[b] = off;
      off += l.bufBytes;
    }
  }
  /* x-side mailboxes */
  const size_t nq = elemSize_.size();
  mboxOff_.assign(27 * nq, 0);
  mboxBytes_ = 0;
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; x += 2) {
        const Dim3 side(x, y, z);
        if (radius_.dir(side) == 0) continue;
        for (size_t qi = 0; qi < nq; ++qi) {
          mboxOff_[size_t(DirectionMap<int>::slot(side)) * nq + qi] = mboxBytes_;
          mboxBytes_ += round_up_multiple(size_t(halo_bytes(side, int64_t(qi))), 256);
        }
      }
  for (int p = 0; p < 2; ++p) {
    off = round_up_multiple(off, slabAlign);
    mboxBase_[p] = off;
    off += mboxBytes_;
  }
  slabBytes_ = round_up_multiple(off, slabAlign);
}

void LocalDomain::realize() {
  STENCIL_RANGE("LocalDomain::realize");
  if (slab_) return;
  DevGuard g(dev_);
  compute_layout();
  LOG_DEBUG("LocalDomain::realize origin=" << origin_ << " size=" << sz_ << " slab=" << slabBytes_ << "B on cuda" << dev_);
  CUDA_RUNTIME(cudaMalloc(&slab_, slabBytes_));
  ownsSlab_ = true;
  /* control block (flags) must start at zero; data is left uninitialised like the reference */
  CUDA_RUNTIME(cudaMemset(slab_, 0, kControlBytes));

  const size_t nq = elemSize_.size();
  if (nq) {
    void **tables = nullptr;
    CUDA_RUNTIME(cudaMalloc(&tables, 2 * nq * sizeof(void *)));
    std::vector<void *> host(2 * nq);
    for (size_t qi = 0; qi < nq; ++qi) {
      host[qi] = buffer(qi, 0);      // table for parity 0: curr = buffer 0
      host[nq + qi] = buffer(qi, 1); // table for parity 1: curr = buffer 1
    }
    CUDA_RUNTIME(cudaMemcpy(tables, host.data(), host.size() * sizeof(void *), cudaMemcpyHostToDevice));
    devPtrTable_[0] = tables;
    devPtrTable_[1] = tables + nq;
    CUDA_RUNTIME(cudaMalloc(&devElemSizes_, nq * sizeof(size_t)));
    CUDA_RUNTIME(cudaMemcpy(devElemSizes_, elemSize_.data(), nq * sizeof(size_t), cudaMemcpyHostToDevice));
  }
  CUDA_RUNTIME(cudaDeviceSynchronize());
}

std::vector<unsigned char> LocalDomain::region_to_host(const Dim3 &pos, const Dim3 &ext, size_t qi) const {
  const size_t es = elemSize_.at(qi);
  std::vector<unsigned char> out(static_cast<size_t>(ext.flatten()) * es);
  if (out.empty()) return out;
  DevGuard g(dev_);
  /* strided D2H straight out of the padded array: no staging buffer, no pack kernel */
  const Dim3 p = pitch(qi);
  cudaMemcpy3DParms parms;
  std::memset(&parms, 0, sizeof(parms));
  parms.srcPtr = make_cudaPitchedPtr(curr_data(qi), size_t(p.x) * es, size_t(p.x) * es, size_t(p.y));
  parms.srcPos = make_cudaPos(size_t(pos.x) * es, size_t(pos.y), size_t(pos.z));
  parms.dstPtr = make_cudaPitchedPtr(out.data(), size_t(ext.x) * es, size_t(ext.x) * es, size_t(ext.y));
  parms.dstPos = make_cudaPos(0, 0, 0);
  parms.extent = make_cudaExtent(size_t(ext.x) * es, size_t(ext.y), size_t(ext.z));
  parms.kind = cudaMemcpyDeviceToHost;
  CUDA_RUNTIME(cudaMemcpy3D(&parms));
  return out;
}

void LocalDomain::region_from_host(const void *host, const Dim3 &pos, const Dim3 &ext, size_t qi) {
  const size_t es = elemSize_.at(qi);
  if (ext.flatten() == 0) return;
  DevGuard g(dev_);
  const Dim3 p = pitch(qi);
  cudaMemcpy3DParms parms;
  std::memset(&parms, 0, sizeof(parms));
  parms.srcPtr = make_cudaPitchedPtr(const_cast<void *>(host), size_t(ext.x) * es, size_t(ext.x) * es, size_t(ext.y));
  parms.srcPos = make_cudaPos(0, 0, 0);
  parms.dstPtr = make_cudaPitchedPtr(curr_data(qi), size_t(p.x) * es, size_t(p.x) * es, size_t(p.y));
  parms.dstPos = make_cudaPos(size_t(pos.x) * es, size_t(pos.y), size_t(pos.z));
  parms.extent = make_cudaExtent(size_t(ext.x) * es, size_t(ext.y), size_t(ext.z));
  parms.kind = cudaMemcpyHostToDevice;
  CUDA_RUNTIME(cudaMemcpy3D(&parms));
  /* a copy from pageable memory may return while its DMA is still in flight on the legacy stream, and the
     library's own streams are non-blocking (they do not wait for the legacy stream): finish it here */
  CUDA_RUNTIME(cudaDeviceSynchronize());
}
