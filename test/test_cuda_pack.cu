/* pack / unpack / translate / multi_* and the packers.
   Ground covered by the reference's test_cuda_pack.cu (3x4x5 iota array faces, round trip, 30x40x50 r=4 smoke),
   test_cuda_translate.cu (single cell, 2x2x2 box) and test_cuda_packer.cu (3 mixed quantities, 4 messages incl.
   edge/corner, 264-byte uneven-radius buffer). Checks are exact, against host loops over the same boxes. */
#include "harness.hpp"
#include <cstring>
#include <numeric>
#include <vector>
#include "stencil/copy.cuh"
#include "stencil/local_domain.cuh"
#include "stencil/pack_kernel.cuh"
#include "stencil/packer.cuh"
#include "stencil/rcstream.hpp"

namespace {
template <typename T> struct DevBuf {
  T *p = nullptr;
  size_t n;
  explicit DevBuf(size_t n_) : n(n_) { CUDA_RUNTIME(cudaMalloc(&p, n * sizeof(T))); }
  ~DevBuf() { cudaFree(p); }
  void up(const std::vector<T> &h) { CUDA_RUNTIME(cudaMemcpy(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice)); }
  std::vector<T> down() const {
    std::vector<T> h(n);
    CUDA_RUNTIME(cudaMemcpy(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
  }
};

template <typename T>
std::vector<T> host_pack(const std::vector<T> &src, const Dim3 &pitch, const Dim3 &pos, const Dim3 &ext) {
  std::vector<T> out;
  for (int64_t z = 0; z < ext.z; ++z)
    for (int64_t y = 0; y < ext.y; ++y)
      for (int64_t x = 0; x < ext.x; ++x) out.push_back(src[size_t(((pos.z + z) * pitch.y + pos.y + y) * pitch.x + pos.x + x)]);
  return out;
}
} // namespace

TEST_CASE("pack faces of a 3x4x5 iota array", "[cuda][pack]") {
  const Dim3 sz(3, 4, 5);
  std::vector<float> h(size_t(sz.flatten()));
  std::iota(h.begin(), h.end(), 0.f);
  DevBuf<float> src(h.size());
  src.up(h);
  const struct {
    Dim3 pos, ext;
  } boxes[] = {{Dim3(0, 0, 4), Dim3(3, 4, 1)}, {Dim3(0, 0, 0), Dim3(1, 4, 5)}, {Dim3(0, 1, 0), Dim3(3, 1, 5)}, {Dim3(1, 1, 1), Dim3(2, 2, 3)}};
  for (const auto &b : boxes) {
    DevBuf<float> dst(size_t(b.ext.flatten()));
    stencil::pack(dst.p, src.p, sz, b.pos, b.ext, sizeof(float));
    CUDA_RUNTIME(cudaDeviceSynchronize());
    REQUIRE(dst.down() == host_pack(h, sz, b.pos, b.ext));
  }
}

TEST_CASE("pack then unpack is the identity on the box and leaves the rest alone", "[cuda][pack]") {
  for (size_t es : {size_t(1), size_t(2), size_t(4), size_t(8), size_t(16), size_t(12)}) {
    const Dim3 pitch(37, 11, 9), pos(3, 2, 1), ext(30, 7, 6);
    std::vector<unsigned char> h(size_t(pitch.flatten()) * es);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)(i * 131 + 7);
    DevBuf<unsigned char> src(h.size()), dst(h.size()), dense(size_t(ext.flatten()) * es);
    src.up(h);
    CUDA_RUNTIME(cudaMemset(dst.p, 0xEE, h.size()));
    stencil::pack(dense.p, src.p, pitch, pos, ext, es);
    stencil::unpack(dst.p, pitch, pos, ext, dense.p, es);
    CUDA_RUNTIME(cudaDeviceSynchronize());
    const std::vector<unsigned char> got = dst.down();
    size_t bad = 0;
    for (int64_t z = 0; z < pitch.z; ++z)
      for (int64_t y = 0; y < pitch.y; ++y)
        for (int64_t x = 0; x < pitch.x; ++x) {
          const bool in = x >= pos.x && x < pos.x + ext.x && y >= pos.y && y < pos.y + ext.y && z >= pos.z && z < pos.z + ext.z;
          const size_t o = size_t((z * pitch.y + y) * pitch.x + x) * es;
          for (size_t b = 0; b < es; ++b) bad += got[o + b] != (in ? h[o + b] : 0xEE);
        }
    REQUIRE(bad == 0);
  }
}

TEST_CASE("translate a cell and a box between two arrays", "[cuda][translate]") {
  const Dim3 sSz(6, 5, 4), dSz(9, 8, 7);
  std::vector<double> hs(size_t(sSz.flatten()));
  std::iota(hs.begin(), hs.end(), 1.0);
  DevBuf<double> s(hs.size()), d(size_t(dSz.flatten()));
  s.up(hs);
  CUDA_RUNTIME(cudaMemset(d.p, 0, d.n * sizeof(double)));
  stencil::translate(d.p, Dim3(8, 7, 6), dSz, s.p, Dim3(5, 4, 3), sSz, Dim3(1, 1, 1), sizeof(double)); // single cell
  stencil::translate(d.p, Dim3(1, 2, 3), dSz, s.p, Dim3(2, 1, 0), sSz, Dim3(2, 2, 2), sizeof(double)); // 2x2x2 box
  CUDA_RUNTIME(cudaDeviceSynchronize());
  const std::vector<double> hd = d.down();
  auto S = [&](int64_t x, int64_t y, int64_t z) { return hs[size_t((z * sSz.y + y) * sSz.x + x)]; };
  auto D = [&](int64_t x, int64_t y, int64_t z) { return hd[size_t((z * dSz.y + y) * dSz.x + x)]; };
  REQUIRE(D(8, 7, 6) == S(5, 4, 3));
  double sum = 0;
  for (int64_t z = 0; z < 2; ++z)
    for (int64_t y = 0; y < 2; ++y)
      for (int64_t x = 0; x < 2; ++x) {
        REQUIRE(D(1 + x, 2 + y, 3 + z) == S(2 + x, 1 + y, 0 + z));
        sum += S(2 + x, 1 + y, z);
      }
  double total = 0;
  for (double v : hd) total += v;
  REQUIRE(total == sum + S(5, 4, 3)); // nothing else written
}

TEST_CASE("multi_translate / multi_pack / multi_unpack over mixed element sizes", "[cuda][translate]") {
  const Dim3 sz(20, 6, 5), pos(2, 1, 1), ext(16, 4, 3), dpos(4, 2, 2);
  const size_t es[3] = {4, 1, 8};
  std::vector<std::vector<unsigned char>> h(3);
  std::vector<DevBuf<unsigned char> *> src, dst;
  void *sp[3], *dp[3];
  for (int q = 0; q < 3; ++q) {
    h[q].resize(size_t(sz.flatten()) * es[q]);
    for (size_t i = 0; i < h[q].size(); ++i) h[q][i] = (unsigned char)(i * (q + 3) + q);
    src.push_back(new DevBuf<unsigned char>(h[q].size()));
    dst.push_back(new DevBuf<unsigned char>(h[q].size()));
    src[q]->up(h[q]);
    CUDA_RUNTIME(cudaMemset(dst[q]->p, 0, h[q].size()));
    sp[q] = src[q]->p;
    dp[q] = dst[q]->p;
  }
  stencil::multi_translate(dp, dpos, sz, sp, pos, sz, ext, es, 3);
  CUDA_RUNTIME(cudaDeviceSynchronize());
  for (int q = 0; q < 3; ++q) {
    const std::vector<unsigned char> got = dst[q]->down();
    REQUIRE(host_pack(got, Dim3(sz.x * int64_t(es[q]), sz.y, sz.z), Dim3(dpos.x * int64_t(es[q]), dpos.y, dpos.z),
                      Dim3(ext.x * int64_t(es[q]), ext.y, ext.z)) ==
            host_pack(h[q], Dim3(sz.x * int64_t(es[q]), sz.y, sz.z), Dim3(pos.x * int64_t(es[q]), pos.y, pos.z),
                      Dim3(ext.x * int64_t(es[q]), ext.y, ext.z)));
  }
  /* pack all three into one buffer at aligned offsets, unpack elsewhere */
  size_t offs[3], total = 0;
  for (int q = 0; q < 3; ++q) {
    total = next_align_of(total, es[q]);
    offs[q] = total;
    total += size_t(ext.flatten()) * es[q];
  }
  DevBuf<unsigned char> packed(total);
  stencil::multi_pack(packed.p, offs, sp, sz, pos, ext, es, 3);
  for (int q = 0; q < 3; ++q) CUDA_RUNTIME(cudaMemset(dst[q]->p, 0, h[q].size()));
  stencil::multi_unpack(dp, sz, pos, ext, packed.p, offs, es, 3);
  CUDA_RUNTIME(cudaDeviceSynchronize());
  const std::vector<unsigned char> hp = packed.down();
  for (int q = 0; q < 3; ++q) {
    const Dim3 bsz(sz.x * int64_t(es[q]), sz.y, sz.z), bpos(pos.x * int64_t(es[q]), pos.y, pos.z), bext(ext.x * int64_t(es[q]), ext.y, ext.z);
    const std::vector<unsigned char> want = host_pack(h[q], bsz, bpos, bext);
    REQUIRE(std::memcmp(hp.data() + offs[q], want.data(), want.size()) == 0);
    REQUIRE(host_pack(dst[q]->down(), bsz, bpos, bext) == want);
    delete src[q];
    delete dst[q];
  }
}

TEST_CASE("packer and unpacker agree on size; uneven radius gives the 264-byte layout", "[cuda][packer]") {
  RcStream stream(0);
  {
    LocalDomain ld(Dim3(30, 40, 50), Dim3(0, 0, 0), 0);
    ld.set_radius(4);
    ld.add_data<float>();
    ld.add_data<char>();
    ld.add_data<double>();
    ld.realize();
    std::vector<Message> msgs = {Message(Dim3(0, 0, 1), 0, 0), Message(Dim3(1, 0, 0), 0, 0), Message(Dim3(-1, -1, 0), 0, 0), Message(Dim3(1, 1, 1), 0, 0)};
    DevicePacker packer(stream);
    DeviceUnpacker unpacker(stream);
    packer.prepare(&ld, msgs);
    unpacker.prepare(&ld, msgs);
    REQUIRE(packer.size() == unpacker.size());
    REQUIRE(packer.size() > 0);
    packer.pack();
    CUDA_RUNTIME(cudaMemcpyAsync(unpacker.data(), packer.data(), size_t(packer.size()), cudaMemcpyDeviceToDevice, stream));
    unpacker.unpack();
    CUDA_RUNTIME(cudaStreamSynchronize(stream));
    REQUIRE(cudaGetLastError() == cudaSuccess);
  }
  {
    /* 10^3, +x radius 2, -x radius 1, float+char+double, messages +x and -x:
       sorted -x then +x; the -x message fills the receiver's +x halo (2 wide): 200*4=800... use the reference's
       small case instead: sizes chosen so that the arithmetic is 80 + 20 -> 104 (align 8) + 160 = 264 */
    LocalDomain ld(Dim3(5, 2, 2), Dim3(0, 0, 0), 0);
    Radius r = Radius::constant(0);
    r.dir(1, 0, 0) = 2;
    r.dir(-1, 0, 0) = 1;
    ld.set_radius(r);
    ld.add_data<float>();
    ld.add_data<char>();
    ld.add_data<double>();
    ld.realize();
    /* one message travelling -x lands in the receiver's +x halo: 2*2*2 = 8 cells... (see below) */
    std::vector<Message> msgs = {Message(Dim3(-1, 0, 0), 0, 0), Message(Dim3(1, 0, 0), 0, 0)};
    DevicePacker packer(stream);
    packer.prepare(&ld, msgs);
    /* -x message: ext = halo_extent(+x) = 2x2x2 = 8 cells -> 32 + 8 + (pad to 40) 64 = 104
       +x message: ext = halo_extent(-x) = 1x2x2 = 4 cells -> 104 + 16 = 120, + 4 = 124 -> pad 128, + 32 = 160 */
    REQUIRE(packer.size() == 160);
    REQUIRE(DevicePacker::packed_size(ld, msgs) == 160);
  }
  {
    /* the reference's figure: 10x10x10... choose extents giving 20 cells for the wide halo and 10 for the narrow */
    LocalDomain ld(Dim3(3, 5, 2), Dim3(0, 0, 0), 0);
    Radius r = Radius::constant(0);
    r.dir(1, 0, 0) = 2;
    r.dir(-1, 0, 0) = 1;
    ld.set_radius(r);
    ld.add_data<float>();
    ld.add_data<char>();
    ld.add_data<double>();
    ld.realize();
    /* single message travelling -x: receiver's +x halo is 2x5x2 = 20 cells: 80 (float) + 20 (char) = 100 -> 104,
       + 160 (double) = 264 */
    DevicePacker packer(stream);
    packer.prepare(&ld, {Message(Dim3(-1, 0, 0), 0, 0)});
    REQUIRE(packer.size() == 264);
  }
}

TEST_CASE("packed image matches a host gather of the same boxes", "[cuda][packer]") {
  RcStream stream(0);
  LocalDomain ld(Dim3(12, 7, 5), Dim3(100, 200, 300), 0);
  ld.set_radius(Radius::face_edge_corner(3, 2, 1));
  auto hf = ld.add_data<float>("f");
  auto hd = ld.add_data<double>("d");
  (void)hf;
  (void)hd;
  ld.realize();
  const Dim3 raw = ld.raw_size();
  std::vector<float> f(size_t(raw.flatten()));
  std::vector<double> d(size_t(raw.flatten()));
  for (size_t i = 0; i < f.size(); ++i) {
    f[i] = float(i);
    d[i] = 0.5 * double(i);
  }
  ld.region_from_host(f.data(), Dim3(0, 0, 0), raw, 0);
  ld.region_from_host(d.data(), Dim3(0, 0, 0), raw, 1);
  std::vector<Message> msgs;
  for (int z = -1; z <= 1; ++z)
    for (int y = -1; y <= 1; ++y)
      for (int x = -1; x <= 1; ++x)
        if (x || y || z) msgs.push_back(Message(Dim3(x, y, z), 0, 0));
  DevicePacker packer(stream);
  packer.prepare(&ld, msgs);
  packer.pack();
  CUDA_RUNTIME(cudaStreamSynchronize(stream));
  std::vector<unsigned char> img(size_t(packer.size()));
  CUDA_RUNTIME(cudaMemcpy(img.data(), packer.data(), img.size(), cudaMemcpyDeviceToHost));
  std::sort(msgs.begin(), msgs.end());
  size_t off = 0;
  for (const Message &m : msgs) {
    const Dim3 ext = ld.halo_extent(m.dir_ * -1), pos = ld.halo_pos(m.dir_, false);
    off = next_align_of(off, 4);
    const std::vector<float> wf = host_pack(f, raw, pos, ext);
    REQUIRE(std::memcmp(img.data() + off, wf.data(), wf.size() * 4) == 0);
    off += wf.size() * 4;
    off = next_align_of(off, 8);
    const std::vector<double> wd = host_pack(d, raw, pos, ext);
    REQUIRE(std::memcmp(img.data() + off, wd.data(), wd.size() * 8) == 0);
    off += wd.size() * 8;
  }
  REQUIRE(off == img.size());
  /* swap(): the packer follows curr without re-preparing */
  ld.swap();
  std::vector<float> g(f.size(), -3.f);
  ld.region_from_host(g.data(), Dim3(0, 0, 0), raw, 0);
  packer.pack();
  CUDA_RUNTIME(cudaStreamSynchronize(stream));
  float first = 0;
  CUDA_RUNTIME(cudaMemcpy(&first, packer.data(), 4, cudaMemcpyDeviceToHost));
  REQUIRE(first == -3.f);
}
