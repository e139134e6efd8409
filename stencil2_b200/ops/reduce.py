"""Deterministic fp64 sum of a box of an fp32 pitched device array (native: src/reduce.cu)."""
import ctypes as C

from .._lib import lib, i3


def reduce_sum_f32(src, pitch, ext, stream=0):
    out = C.c_double()
    rc = lib().stencil_reduce_sum_f32(src, i3(pitch), i3(ext), stream or None, C.byref(out))
    if rc:
        raise ValueError("stencil_reduce_sum_f32 failed")
    return out.value
