"""RankPartition / NodePartition through the native library (reference: include/stencil/partition.hpp:23-310)."""
from .._lib import lib, i3, radius27, I64x3


def set_partition_ties(x_first):
    """Axis cut on ties: False = z, y, x (default: long x rows, contiguous faces); True = x first (reference order).
    Returns the previous setting."""
    return bool(lib().stencil_set_partition_ties(1 if x_first else 0))


class RankPartition:
    """Flat split of `size` into n boxes: prime factors of n largest first, always cutting the longest axis."""

    def __init__(self, size, n):
        self.size, self.n = tuple(size), int(n)
        d = I64x3()
        if lib().stencil_rank_partition(i3(size), self.n, d) != 0:
            raise ValueError("bad partition arguments")
        self.dim = tuple(d)

    def subdomain(self, idx):
        s, o = I64x3(), I64x3()
        if lib().stencil_rank_partition_subdomain(i3(self.size), self.n, i3(idx), s, o) != 0:
            raise IndexError(idx)
        return tuple(s), tuple(o)


class NodePartition:
    """Two-level split (nodes, then GPUs per node) along the axis with the smallest radius-weighted interface."""

    def __init__(self, size, radius, nodes, gpus):
        self.size, self.radius, self.nodes, self.gpus = tuple(size), radius, int(nodes), int(gpus)
        s, n = I64x3(), I64x3()
        if lib().stencil_node_partition(i3(size), radius27(radius), self.nodes, self.gpus, s, n) != 0:
            raise ValueError("bad partition arguments")
        self.sys_dim, self.node_dim = tuple(s), tuple(n)
        self.dim = tuple(a * b for a, b in zip(self.sys_dim, self.node_dim))

    def subdomain(self, idx):
        s, o = I64x3(), I64x3()
        if lib().stencil_node_partition_subdomain(i3(self.size), radius27(self.radius), self.nodes, self.gpus, i3(idx), s, o) != 0:
            raise IndexError(idx)
        return tuple(s), tuple(o)


def rank_partition(size, n):
    return RankPartition(size, n)


def node_partition(size, radius, nodes, gpus):
    return NodePartition(size, radius, nodes, gpus)
