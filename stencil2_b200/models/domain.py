"""DistributedDomain: configure -> realize() -> { compute; exchange(); swap() } (reference: include/stencil/stencil.hpp)."""
import ctypes as C

import numpy as np

from .._lib import lib, i3, radius27, I64x3


class Method:
    NONE = 0
    CUDA_MPI = 1          # host-staged
    CUDA_AWARE_MPI = 2
    COLOCATED = 4         # CUDA IPC + cudaMemcpyPeer between processes of a node
    MEMCPY_PEER = 8       # pack + cudaMemcpyPeer inside a process
    KERNEL = 16           # same-GPU copy kernel
    NVLINK_DIRECT = 32    # device-initiated stores into the neighbour's halo (new)
    ALL = 1 + 4 + 8 + 16 + 32


class Placement:
    NODE_AWARE = 0
    TRIVIAL = 1


class _Sub:
    """One local subdomain: geometry plus host transfers of its quantities."""

    def __init__(self, dd, di):
        self._dd, self.index = dd, di
        o, s, r = I64x3(), I64x3(), I64x3()
        gpu = C.c_int()
        lib().stencil_subdomain_info(dd._h, di, o, s, r, C.byref(gpu))
        self.origin, self.size, self.raw_size, self.gpu = tuple(o), tuple(s), tuple(r), gpu.value

    def quantity(self, qi):
        """(pitch in elements, curr device pointer, next device pointer, element size)"""
        p = I64x3()
        cur, nxt, es = C.c_void_p(), C.c_void_p(), C.c_size_t()
        if lib().stencil_subdomain_quantity(self._dd._h, self.index, qi, p, C.byref(cur), C.byref(nxt), C.byref(es)) != 0:
            raise IndexError(qi)
        return tuple(p), cur.value, nxt.value, es.value

    def to_host(self, qi, dtype, pos=None, ext=None):
        """Dense numpy array [z, y, x] of a box of the allocation (default: everything incl. halos)."""
        pos = (0, 0, 0) if pos is None else pos
        ext = self.raw_size if ext is None else ext
        out = np.empty((ext[2], ext[1], ext[0]), dtype=dtype)
        rc = lib().stencil_subdomain_to_host(self._dd._h, self.index, qi, i3(pos), i3(ext), out.ctypes.data)
        if rc:
            raise ValueError("to_host failed")
        return out

    def from_host(self, qi, array, pos=(0, 0, 0)):
        a = np.ascontiguousarray(array)
        ext = (a.shape[2], a.shape[1], a.shape[0])
        rc = lib().stencil_subdomain_from_host(self._dd._h, self.index, qi, i3(pos), i3(ext), a.ctypes.data)
        if rc:
            raise ValueError("from_host failed")

    def interior(self):
        lo, hi = I64x3(), I64x3()
        lib().stencil_subdomain_interior(self._dd._h, self.index, lo, hi)
        return tuple(lo), tuple(hi)

    def exterior(self):
        lo = ((C.c_int64 * 3) * 6)()
        hi = ((C.c_int64 * 3) * 6)()
        n = lib().stencil_subdomain_exterior(self._dd._h, self.index, lo, hi)
        return [(tuple(lo[k]), tuple(hi[k])) for k in range(n)]


class DistributedDomain:
    def __init__(self, x, y, z, _borrowed=None):
        self._owned = _borrowed is None
        self._h = lib().stencil_domain_create(x, y, z) if self._owned else _borrowed
        self._realized = not self._owned
        self.size = (x, y, z)

    def close(self):
        if self._h and self._owned:
            lib().stencil_domain_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_radius(self, r):
        lib().stencil_domain_set_radius(self._h, radius27(r))

    def add_data(self, dtype, name=""):
        return lib().stencil_domain_add_data(self._h, np.dtype(dtype).itemsize, name.encode())

    def set_methods(self, methods):
        lib().stencil_domain_set_methods(self._h, int(methods))

    def set_placement(self, placement):
        lib().stencil_domain_set_placement(self._h, int(placement))

    def set_gpus(self, ids):
        arr = (C.c_int * len(ids))(*ids)
        lib().stencil_domain_set_gpus(self._h, arr, len(ids))

    use_gpus = set_gpus

    def set_plan_file(self, on):
        lib().stencil_domain_set_plan_file(self._h, 1 if on else 0)

    def realize(self):
        lib().stencil_domain_realize(self._h)
        self._realized = True

    def exchange(self):
        lib().stencil_domain_exchange(self._h)

    def exchange_async(self):
        """Enqueue the device work of an exchange and return; pair with exchange_wait()."""
        lib().stencil_domain_exchange_async(self._h)

    def exchange_wait(self):
        lib().stencil_domain_exchange_wait(self._h)

    @property
    def exchange_device_only(self):
        """True if no message of the plan needs the host (then torch streams can be chained to the exchange by events)."""
        return lib().stencil_domain_exchange_device_only(self._h) == 1

    def exchange_done_event(self, subdomain=0):
        """Raw cudaEvent_t (int) recorded at the end of the last exchange_async() on that subdomain's exchange stream,
        e.g. for torch.cuda.ExternalStream / cudaStreamWaitEvent through cuda-python. None if the plan is not device-only."""
        return lib().stencil_domain_exchange_done_event(self._h, int(subdomain))

    def exchange_after(self, subdomain, cuda_event):
        """The next exchange_async() starts after `cuda_event` (raw cudaEvent_t handle, e.g. torch.cuda.Event().cuda_event)."""
        lib().stencil_domain_exchange_after(self._h, int(subdomain), cuda_event)

    def time_exchange(self, iters):
        """(host seconds, device seconds) of `iters` blocking exchanges, max over ranks (what weak-exchange prints)."""
        dev = C.c_double(0)
        t = lib().stencil_domain_time_exchange(self._h, int(iters), C.byref(dev))
        return t, dev.value

    @property
    def grid(self):
        """Subdomains along (x, y, z) over the whole job."""
        g = I64x3()
        lib().stencil_domain_grid(self._h, g)
        return tuple(g)

    def swap(self):
        lib().stencil_domain_swap(self._h)

    def exchange_bytes_for_method(self, methods):
        return lib().stencil_domain_exchange_bytes(self._h, int(methods))

    def write_paraview(self, prefix, zero_nans=False):
        lib().stencil_domain_write_paraview(self._h, prefix.encode(), 1 if zero_nans else 0)

    @property
    def subdomains(self):
        return [_Sub(self, di) for di in range(lib().stencil_domain_num_subdomains(self._h))]
