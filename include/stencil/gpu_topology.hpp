#pragma once
/* gpu_topo: GPU-GPU "bandwidth" model for placement + registry of enabled peer access.
   Parity: reference include/stencil/gpu_topology.hpp:7-16, src/gpu_topology.cpp:13-139 (distance 0.1 same GPU,
   1.0 NVLink, 2..7 by PCIe common ancestor; bandwidth = 1/distance; enable_peer/peer).

   B200 difference: on an NVSwitch machine no NVLink of a GPU ends at another GPU — all 18 end at
   switches — so "is the remote PCI device my peer?" is always false. Here a pair counts as NVLink-connected when
   nvmlDeviceGetP2PStatus(NVLINK) says OK, or both GPUs have active links whose remote device type is a switch.
   A measured matrix (from measure-buf-exchange) or a synthetic one can be installed with set_bandwidth_matrix(). */
#include <vector>

#include "stencil/mat2d.hpp"

namespace gpu_topo {

/* relative bandwidth between two CUDA devices of this node (larger = faster); CUDA ordinal ids */
double bandwidth(int src, int dst);
/* override the model: m[i][j] for CUDA ids i,j (empty matrix restores the NVML/CUDA probe) */
void set_bandwidth_matrix(const Mat2D<double> &m);
/* the matrix the probe (or override) yields for devices 0..n-1 */
Mat2D<double> bandwidth_matrix(int n);

/* enable src->dst peer access if the hardware allows; remembers the outcome */
void enable_peer(int src, int dst);
/* Record whether src could reach dst without enabling anything: for a dst that belongs to ANOTHER process.
   Its memory is only ever reached through cudaIpcOpenMemHandle(..., cudaIpcMemLazyEnablePeerAccess), which sets up
   what it needs; cudaDeviceEnablePeerAccess() here would create a context on that GPU in this process too (with one
   process per GPU: 64 contexts on an 8-GPU box, and 13 s of the 14 s realize() took). */
void note_peer(int src, int dst);
/* was peer access from src to dst enabled by enable_peer() (true for src == dst once registered) */
bool peer(int src, int dst);
/* true if every active NVLink of `dev` ends at an NVSwitch */
bool behind_nvswitch(int dev);

} // namespace gpu_topo
