#pragma once
/* NVML access. Parity: reference include/stencil/nvml.hpp:7-25 (NVML() check macro + nvml::lazy_init()).
   Difference: libnvidia-ml is dlopen()ed at first use instead of linked, so the library also loads on machines
   with no NVIDIA driver (CPU-only CI); there nvml::available() is false and topology falls back to CUDA queries. */
#include <cstddef>

namespace nvml {

/* minimal mirror of the NVML ABI we touch (stable since NVML 10) */
typedef struct nvmlDevice_st *Device;
struct PciInfo {
  char busIdLegacy[16];
  unsigned int domain, bus, device, pciDeviceId, pciSubSystemId;
  char busId[32];
};
enum { SUCCESS = 0 };
enum TopologyLevel { TOPO_INTERNAL = 0, TOPO_SINGLE = 10, TOPO_MULTIPLE = 20, TOPO_HOSTBRIDGE = 30, TOPO_NODE = 40, TOPO_SYSTEM = 50 };
enum RemoteDeviceType { REMOTE_GPU = 0, REMOTE_IBMNPU = 1, REMOTE_SWITCH = 2, REMOTE_UNKNOWN = 255 };
enum P2PIndex { P2P_READ = 0, P2P_WRITE = 1, P2P_NVLINK = 2, P2P_ATOMICS = 3 };

struct Api {
  int (*init)();
  const char *(*errorString)(int);
  int (*deviceGetCount)(unsigned int *);
  int (*deviceGetHandleByIndex)(unsigned int, Device *);
  int (*deviceGetHandleByPciBusId)(const char *, Device *);
  int (*deviceGetPciInfo)(Device, PciInfo *);
  int (*deviceGetNvLinkState)(Device, unsigned int, unsigned int *);
  int (*deviceGetNvLinkRemotePciInfo)(Device, unsigned int, PciInfo *);
  int (*deviceGetNvLinkRemoteDeviceType)(Device, unsigned int, int *);
  int (*deviceGetTopologyCommonAncestor)(Device, Device, int *);
  int (*deviceGetP2PStatus)(Device, Device, int, int *);
};

/* load + nvmlInit once; returns false when there is no usable NVML */
bool lazy_init();
bool available();
const Api &api();

} // namespace nvml

#define NVML(stmt)                                                                                                     \
  {                                                                                                                    \
    int nvml_rc_ = (stmt);                                                                                             \
    if (nvml_rc_ != ::nvml::SUCCESS) {                                                                                 \
      std::fprintf(stderr, "%s:%d: NVML error: %s\n", __FILE__, __LINE__, ::nvml::api().errorString(nvml_rc_));       \
      std::exit(-1);                                                                                                   \
    }                                                                                                                  \
  }
