"""AstarothSim: the reference's multi-field radius-3 workload (bin/astaroth_sim.cu) on the native engine
(src/astaroth.cu): `fields` quantities of fp32 or fp64, the reference's 7-point stand-in, a real 19-point
6th-order Laplacian step, or that step plus cross-derivative terms (reads all 26 halo directions); TMA-pipelined
kernel on B200."""
import ctypes as C

import numpy as np

from .._lib import lib
from .domain import DistributedDomain, Method, Placement


class AstarothSim:
    AVERAGE7, STAR19, STAR19_CROSS = 0, 1, 2

    def __init__(self, x, y, z, fields=1, fp64=False, stencil=0, overlap=None, methods=Method.ALL, placement=Placement.NODE_AWARE,
                 gpus=None, dt=0.01):
        """overlap: True = interior / exchange / exterior (reference schedule), False = exchange then one launch,
        None = let the engine choose."""
        arr = (C.c_int * len(gpus))(*gpus) if gpus else None
        ov = -1 if overlap is None else (1 if overlap else 0)
        self._h = lib().stencil_astaroth_create(x, y, z, int(methods), int(placement), int(stencil), int(fields), 1 if fp64 else 0, ov, arr,
                                                len(gpus) if gpus else 0, float(dt))
        if not self._h:
            raise ValueError("bad AstarothSim arguments")
        self.size, self.fields, self.fp64, self.stencil, self.dt = (x, y, z), int(fields), bool(fp64), int(stencil), float(dt)
        self.domain = DistributedDomain(x, y, z, _borrowed=lib().stencil_astaroth_domain(self._h))

    def close(self):
        if self._h:
            lib().stencil_astaroth_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def init(self):
        lib().stencil_astaroth_init(self._h)

    def step(self):
        lib().stencil_astaroth_step(self._h)

    def sync(self):
        lib().stencil_astaroth_sync(self._h)

    def run(self, iters):
        """(seconds, seconds inside exchange()) for `iters` iterations, host wall clock, max over ranks."""
        ex = C.c_double(0)
        t = lib().stencil_astaroth_run(self._h, int(iters), C.byref(ex))
        return t, ex.value

    def run_timed(self, iters):
        """(host seconds, seconds inside exchange(), device seconds between CUDA events around the loop)."""
        ex, dev = C.c_double(0), C.c_double(0)
        t = lib().stencil_astaroth_run_timed(self._h, int(iters), C.byref(ex), C.byref(dev))
        return t, ex.value, dev.value

    @property
    def launches(self):
        return lib().stencil_astaroth_launches(self._h)

    def gather(self, field):
        """Dense float64 array [z, y, x] of one field (NaN where another rank owns the cell)."""
        x, y, z = self.size
        out = np.empty((z, y, x), dtype=np.float64)
        rc = lib().stencil_astaroth_gather(self._h, int(field), out.ctypes.data_as(C.c_void_p), out.size)
        if rc != 0:
            raise ValueError(f"gather failed ({rc})")
        return out


def astaroth_host_step(curr, stencil=1, fp64=True, dt=0.01):
    """Host reference of one step on a dense periodic [z, y, x] float64 array (same arithmetic order as the kernel)."""
    curr = np.ascontiguousarray(curr, dtype=np.float64)
    z, y, x = curr.shape
    nxt = np.empty_like(curr)
    rc = lib().stencil_astaroth_host_step(nxt.ctypes.data_as(C.c_void_p), curr.ctypes.data_as(C.c_void_p), x, y, z, int(stencil),
                                          1 if fp64 else 0, float(dt))
    if rc != 0:
        raise ValueError("bad arguments")
    return nxt
