/* GPU topology probe and peer-access registry. Reference: src/gpu_topology.cpp:13-139. */
#include "stencil/gpu_topology.hpp"

#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>

#include "stencil/cuda_runtime.hpp"
#include "stencil/logging.hpp"
#include "stencil/nvml.hpp"

/* ---------------- NVML through dlopen ---------------- */
namespace nvml {
namespace {
Api g_api{};
bool g_ok = false;
std::once_flag g_once;

template <typename F> bool bind(void *lib, F &fn, const char *a, const char *b = nullptr) {
  void *p = dlsym(lib, a);
  if (!p && b) p = dlsym(lib, b);
  fn = reinterpret_cast<F>(p);
  return p != nullptr;
}

void load() {
  void *lib = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
  if (!lib) lib = dlopen("libnvidia-ml.so", RTLD_NOW | RTLD_LOCAL);
  if (!lib) return;
  bool ok = true;
  ok &= bind(lib, g_api.init, "nvmlInit_v2", "nvmlInit");
  ok &= bind(lib, g_api.errorString, "nvmlErrorString");
  ok &= bind(lib, g_api.deviceGetCount, "nvmlDeviceGetCount_v2", "nvmlDeviceGetCount");
  ok &= bind(lib, g_api.deviceGetHandleByIndex, "nvmlDeviceGetHandleByIndex_v2", "nvmlDeviceGetHandleByIndex");
  ok &= bind(lib, g_api.deviceGetHandleByPciBusId, "nvmlDeviceGetHandleByPciBusId_v2", "nvmlDeviceGetHandleByPciBusId");
  ok &= bind(lib, g_api.deviceGetPciInfo, "nvmlDeviceGetPciInfo_v3", "nvmlDeviceGetPciInfo_v2");
  ok &= bind(lib, g_api.deviceGetTopologyCommonAncestor, "nvmlDeviceGetTopologyCommonAncestor");
  /* optional entry points */
  bind(lib, g_api.deviceGetNvLinkState, "nvmlDeviceGetNvLinkState");
  bind(lib, g_api.deviceGetNvLinkRemotePciInfo, "nvmlDeviceGetNvLinkRemotePciInfo_v2", "nvmlDeviceGetNvLinkRemotePciInfo");
  bind(lib, g_api.deviceGetNvLinkRemoteDeviceType, "nvmlDeviceGetNvLinkRemoteDeviceType");
  bind(lib, g_api.deviceGetP2PStatus, "nvmlDeviceGetP2PStatus");
  if (!ok) return;
  g_ok = g_api.init() == SUCCESS;
}
} // namespace

bool lazy_init() {
  std::call_once(g_once, load);
  return g_ok;
}
bool available() { return lazy_init(); }
const Api &api() {
  lazy_init();
  return g_api;
}
} // namespace nvml

/* ---------------- gpu_topo ---------------- */
namespace gpu_topo {
namespace {

std::map<std::pair<int, int>, bool> g_peer;
Mat2D<double> g_override;
std::mutex g_mu;

constexpr double kSame = 0.1, kNvlink = 1.0;
constexpr int kMaxLinks = 18; // NVLink 5: 18 links per GPU

/* NVML handle of a CUDA ordinal, matched by PCI bus id so CUDA_VISIBLE_DEVICES reordering cannot confuse us */
bool nvml_handle(int cudaDev, nvml::Device *out) {
  char busId[32] = {0};
  if (cudaDeviceGetPCIBusId(busId, sizeof(busId), cudaDev) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return nvml::api().deviceGetHandleByPciBusId(busId, out) == nvml::SUCCESS;
}

double pcie_distance(nvml::Device a, nvml::Device b) {
  int level = nvml::TOPO_SYSTEM;
  if (nvml::api().deviceGetTopologyCommonAncestor(a, b, &level) != nvml::SUCCESS) return 7.0;
  switch (level) {
  case nvml::TOPO_INTERNAL: return 2.0;
  case nvml::TOPO_SINGLE: return 3.0;
  case nvml::TOPO_MULTIPLE: return 4.0;
  case nvml::TOPO_HOSTBRIDGE: return 5.0;
  case nvml::TOPO_NODE: return 6.0;
  default: return 7.0;
  }
}

/* 0 = no active link, 1 = links to GPUs only / mixed, 2 = all active links end at switches */
int link_class(nvml::Device d, nvml::Device peer, bool *directToPeer) {
  const nvml::Api &n = nvml::api();
  *directToPeer = false;
  if (!n.deviceGetNvLinkState) return 0;
  int active = 0, toSwitch = 0;
  for (int l = 0; l < kMaxLinks; ++l) {
    unsigned st = 0;
    if (n.deviceGetNvLinkState(d, unsigned(l), &st) != nvml::SUCCESS || !st) continue;
    ++active;
    int type = nvml::REMOTE_UNKNOWN;
    if (n.deviceGetNvLinkRemoteDeviceType && n.deviceGetNvLinkRemoteDeviceType(d, unsigned(l), &type) == nvml::SUCCESS &&
        type == nvml::REMOTE_SWITCH) {
      ++toSwitch;
      continue;
    }
    nvml::PciInfo pci;
    nvml::Device remote;
    if (n.deviceGetNvLinkRemotePciInfo && n.deviceGetNvLinkRemotePciInfo(d, unsigned(l), &pci) == nvml::SUCCESS &&
        n.deviceGetHandleByPciBusId(pci.busId, &remote) == nvml::SUCCESS && remote == peer)
      *directToPeer = true;
  }
  if (!active) return 0;
  return toSwitch == active ? 2 : 1;
}

double probe_distance(int src, int dst) {
  if (src == dst) return kSame;
  if (nvml::available()) {
    nvml::Device a, b;
    if (nvml_handle(src, &a) && nvml_handle(dst, &b)) {
      int status = -1;
      if (nvml::api().deviceGetP2PStatus && nvml::api().deviceGetP2PStatus(a, b, nvml::P2P_NVLINK, &status) == nvml::SUCCESS &&
          status == 0)
        return kNvlink;
      bool direct = false, dummy = false;
      const int ca = link_class(a, b, &direct);
      if (direct) return kNvlink;
      if (ca == 2 && link_class(b, a, &dummy) == 2) return kNvlink; // both hang off the NVSwitch fabric
      return pcie_distance(a, b);
    }
  }
  /* no NVML: peer-capable pairs are "near", others "far" */
  int can = 0;
  if (cudaDeviceCanAccessPeer(&can, src, dst) != cudaSuccess) (void)cudaGetLastError();
  return can ? kNvlink : 6.0;
}

} // namespace

double bandwidth(int src, int dst) {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_override.shape().x > 0) {
      if (src >= 0 && dst >= 0 && uint64_t(src) < g_override.shape().y && uint64_t(dst) < g_override.shape().x)
        return g_override.at(src, dst);
      return src == dst ? 1.0 / kSame : 1.0 / 7.0;
    }
  }
  return 1.0 / probe_distance(src, dst);
}

void set_bandwidth_matrix(const Mat2D<double> &m) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_override = m;
}

Mat2D<double> bandwidth_matrix(int n) {
  Mat2D<double> m(n, n, 0.0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) m.at(i, j) = bandwidth(i, j);
  return m;
}

bool behind_nvswitch(int dev) {
  if (!nvml::available()) return false;
  nvml::Device d;
  if (!nvml_handle(dev, &d)) return false;
  bool direct = false;
  return link_class(d, d, &direct) == 2;
}

void enable_peer(int src, int dst) {
  bool ok = false;
  if (src == dst) {
    ok = true;
  } else {
    int can = 0;
    CUDA_RUNTIME(cudaDeviceCanAccessPeer(&can, src, dst));
    if (can) {
      int prev = 0;
      CUDA_RUNTIME(cudaGetDevice(&prev));
      CUDA_RUNTIME(cudaSetDevice(src));
      const cudaError_t e = cudaDeviceEnablePeerAccess(dst, 0);
      (void)cudaGetLastError();
      CUDA_RUNTIME(cudaSetDevice(prev));
      if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled)
        ok = true;
      else if (e == cudaErrorInvalidDevice)
        LOG_WARN(src << " -> " << dst << " peer access: invalid device");
      else
        CUDA_RUNTIME(e);
    }
  }
  LOG_DEBUG(src << " -> " << dst << " peer access " << (ok ? "enabled" : "unavailable"));
  std::lock_guard<std::mutex> lk(g_mu);
  g_peer[{src, dst}] = ok;
}

void note_peer(int src, int dst) {
  int can = src == dst ? 1 : 0;
  if (src != dst) CUDA_RUNTIME(cudaDeviceCanAccessPeer(&can, src, dst));
  LOG_DEBUG(src << " -> " << dst << " peer access " << (can ? "possible (enabled lazily by the IPC mapping)" : "unavailable"));
  std::lock_guard<std::mutex> lk(g_mu);
  g_peer[{src, dst}] = can != 0;
}

bool peer(int src, int dst) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_peer.find({src, dst});
  return it != g_peer.end() && it->second;
}

} // namespace gpu_topo
