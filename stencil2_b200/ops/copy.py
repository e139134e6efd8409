"""Box copies between pitched device arrays (reference kernels K1-K7; native: src/box_copy.cu). Pointers are plain
ints (e.g. ``tensor.data_ptr()``); pitches/positions/extents are (x, y, z) in elements."""
from .._lib import lib, i3


def pack(dst, src, src_pitch, src_pos, ext, elem_size, stream=0):
    rc = lib().stencil_pack(dst, src, i3(src_pitch), i3(src_pos), i3(ext), elem_size, stream or None)
    if rc:
        raise ValueError("stencil_pack failed")


def unpack(dst, dst_pitch, dst_pos, ext, src, elem_size, stream=0):
    rc = lib().stencil_unpack(dst, i3(dst_pitch), i3(dst_pos), i3(ext), src, elem_size, stream or None)
    if rc:
        raise ValueError("stencil_unpack failed")


def translate(dst, dst_pos, dst_pitch, src, src_pos, src_pitch, ext, elem_size, stream=0):
    rc = lib().stencil_translate(dst, i3(dst_pos), i3(dst_pitch), src, i3(src_pos), i3(src_pitch), i3(ext), elem_size, stream or None)
    if rc:
        raise ValueError("stencil_translate failed")
