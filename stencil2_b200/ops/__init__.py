"""Standalone device operations on raw device pointers (ints): pack / unpack / translate / reduce."""
from .copy import pack, unpack, translate  # noqa: F401
from .reduce import reduce_sum_f32  # noqa: F401
