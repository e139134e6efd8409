"""Jacobi3D: the reference's 7-point heat solver (bin/jacobi3d.cu) on the fused compute+push kernel."""
import ctypes as C

import numpy as np

from .._lib import lib
from .domain import DistributedDomain, Method, Placement


class Jacobi3D:
    FUSED, OVERLAP, NO_OVERLAP = 0, 1, 2

    def __init__(self, x, y, z, methods=Method.ALL, placement=Placement.NODE_AWARE, mode=0, gpus=None, tile_z=0):
        arr = (C.c_int * len(gpus))(*gpus) if gpus else None
        self._h = lib().stencil_jacobi_create(x, y, z, int(methods), int(placement), int(mode), arr, len(gpus) if gpus else 0, tile_z)
        self.size = (x, y, z)
        self.domain = DistributedDomain(x, y, z, _borrowed=lib().stencil_jacobi_domain(self._h))

    def close(self):
        if self._h:
            lib().stencil_jacobi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def mode(self):
        return lib().stencil_jacobi_mode(self._h)

    def init(self):
        lib().stencil_jacobi_init(self._h)

    def init_random(self, seed=1):
        """Initial field = pseudo-random values in [0,1) that depend only on the seed and the global cell coordinate."""
        lib().stencil_jacobi_init_random(self._h, int(seed))

    def fingerprint(self):
        """64-bit fingerprint of the whole field over all ranks; equal between two runs exactly when every cell agrees."""
        return int(lib().stencil_jacobi_fingerprint(self._h))

    def step(self):
        lib().stencil_jacobi_step(self._h)

    def step_io(self, params_ptr, nbytes, result_ptr):
        """One step with a pinned-host parameter upload before it and a pinned-host result download after it."""
        if lib().stencil_jacobi_step_io(self._h, params_ptr, nbytes, result_ptr) != 0:
            raise ValueError("step_io failed")

    def sync(self):
        lib().stencil_jacobi_sync(self._h)

    def run(self, iters):
        """Device-timed seconds for `iters` iterations (max over local GPUs and ranks)."""
        return lib().stencil_jacobi_run(self._h, int(iters))

    def checksum(self):
        return lib().stencil_jacobi_checksum(self._h)

    def probe(self, x, y, z):
        v = C.c_float()
        return v.value if lib().stencil_jacobi_probe(self._h, x, y, z, C.byref(v)) == 1 else None

    @property
    def launches(self):
        return lib().stencil_jacobi_launches(self._h)

    def gather(self):
        """The compute regions of this rank assembled into a dense [z, y, x] array of the global size (cells owned
        by other ranks are NaN)."""
        X, Y, Z = self.size
        out = np.full((Z, Y, X), np.nan, dtype=np.float32)
        self.sync()
        for s in self.domain.subdomains:
            lo = tuple(o for o in s.origin)
            r = [(a - b) // 2 for a, b in zip(s.raw_size, s.size)]  # symmetric halos here (radius 1 faces)
            blk = s.to_host(0, np.float32, pos=(1, 1, 1), ext=s.size)
            del r
            out[lo[2]:lo[2] + s.size[2], lo[1]:lo[1] + s.size[1], lo[0]:lo[0] + s.size[0]] = blk
        return out


def jacobi_host_step(curr):
    """One iteration on a dense periodic [z, y, x] fp32 array with the kernel's exact arithmetic (native host loop)."""
    a = np.ascontiguousarray(curr, dtype=np.float32)
    out = np.empty_like(a)
    Z, Y, X = a.shape
    lib().stencil_jacobi_host_step(out.ctypes.data, a.ctypes.data, X, Y, Z)
    return out
