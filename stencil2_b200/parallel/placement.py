"""QAP solvers and node-aware placement (reference: include/stencil/qap.hpp, partition.hpp:573-864)."""
import ctypes as C

from .._lib import lib, i3, radius27, I64x3


def qap_solve(weight, distance, exact=True):
    """Minimise sum_ab w[a][b] * d[f[a]][f[b]] (0*inf = 0). Returns (f, cost)."""
    n = len(weight)
    W = (C.c_double * (n * n))(*[float(v) for row in weight for v in row])
    D = (C.c_double * (n * n))(*[float(v) for row in distance for v in row])
    f = (C.c_uint64 * n)()
    cost = C.c_double()
    if lib().stencil_qap_solve(W, D, n, 1 if exact else 0, f, C.byref(cost)) != 0:
        raise ValueError("bad QAP arguments")
    return list(f), cost.value


def place_node_aware(size, radius, bandwidth):
    """Place len(bandwidth) subdomains of one node on its GPUs. Returns (cuda id per subdomain, subdomain grid)."""
    n = len(bandwidth)
    B = (C.c_double * (n * n))(*[float(v) for row in bandwidth for v in row])
    cuda = (C.c_int * n)()
    dim = I64x3()
    if lib().stencil_place_node_aware(i3(size), radius27(radius), n, B, cuda, dim) != 0:
        raise ValueError("bad placement arguments")
    return list(cuda), tuple(dim)
