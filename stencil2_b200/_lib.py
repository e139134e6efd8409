"""ctypes loader and prototypes for libstencil_b200.so (include/stencil_c.h)."""
import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "lib", "libstencil_b200.so")
_lock = threading.Lock()
_lib = None


class LibraryNotBuilt(RuntimeError):
    pass


def library_path() -> str:
    return _LIB_PATH


I64x3 = C.c_int64 * 3
U64x27 = C.c_uint64 * 27


def _declare(l):
    vp, i, u64, i64, dbl, sz = C.c_void_p, C.c_int, C.c_uint64, C.c_int64, C.c_double, C.c_size_t
    P = C.POINTER
    sig = {
        "stencil_version": (C.c_char_p, []),
        "stencil_device_count": (i, []),
        "stencil_host_alloc_pinned": (vp, [sz]),
        "stencil_host_free_pinned": (None, [vp]),
        "stencil_comm_init_tcp": (i, [i, i, C.c_char_p, i]),
        "stencil_comm_rank": (i, []),
        "stencil_comm_size": (i, []),
        "stencil_comm_barrier": (None, []),
        "stencil_comm_allreduce_max": (dbl, [dbl]),
        "stencil_comm_finalize": (None, []),
        "stencil_set_partition_ties": (i, [i]),
        "stencil_rank_partition": (i, [I64x3, i64, I64x3]),
        "stencil_rank_partition_subdomain": (i, [I64x3, i64, I64x3, I64x3, I64x3]),
        "stencil_node_partition": (i, [I64x3, U64x27, i64, i64, I64x3, I64x3]),
        "stencil_node_partition_subdomain": (i, [I64x3, U64x27, i64, i64, I64x3, I64x3, I64x3]),
        "stencil_qap_solve": (i, [P(dbl), P(dbl), i, i, P(u64), P(dbl)]),
        "stencil_place_node_aware": (i, [I64x3, U64x27, i, P(dbl), P(i), I64x3]),
        "stencil_gpu_bandwidth": (dbl, [i, i]),
        "stencil_gpu_enable_peer": (i, [i, i]),
        "stencil_pack": (i, [vp, vp, I64x3, I64x3, I64x3, sz, vp]),
        "stencil_unpack": (i, [vp, I64x3, I64x3, I64x3, vp, sz, vp]),
        "stencil_translate": (i, [vp, I64x3, I64x3, vp, I64x3, I64x3, I64x3, sz, vp]),
        "stencil_reduce_sum_f32": (i, [vp, I64x3, I64x3, vp, P(dbl)]),
        "stencil_reduce_sum_f32_noalloc": (i, [vp, I64x3, I64x3, vp, vp]),
        "stencil_domain_create": (vp, [u64, u64, u64]),
        "stencil_domain_destroy": (None, [vp]),
        "stencil_domain_set_radius": (i, [vp, U64x27]),
        "stencil_domain_set_radius_constant": (i, [vp, u64]),
        "stencil_domain_add_data": (i, [vp, sz, C.c_char_p]),
        "stencil_domain_set_methods": (i, [vp, i]),
        "stencil_domain_set_placement": (i, [vp, i]),
        "stencil_domain_set_gpus": (i, [vp, P(i), i]),
        "stencil_domain_set_plan_file": (i, [vp, i]),
        "stencil_domain_realize": (i, [vp]),
        "stencil_domain_exchange": (i, [vp]),
        "stencil_domain_exchange_async": (i, [vp]),
        "stencil_domain_exchange_wait": (i, [vp]),
        "stencil_domain_exchange_device_only": (i, [vp]),
        "stencil_domain_exchange_done_event": (vp, [vp, i]),
        "stencil_domain_exchange_after": (i, [vp, i, vp]),
        "stencil_domain_swap": (i, [vp]),
        "stencil_domain_num_subdomains": (i, [vp]),
        "stencil_domain_grid": (i, [vp, I64x3]),
        "stencil_domain_time_exchange": (dbl, [vp, i, P(dbl)]),
        "stencil_domain_exchange_bytes": (u64, [vp, i]),
        "stencil_domain_write_paraview": (i, [vp, C.c_char_p, i]),
        "stencil_subdomain_info": (i, [vp, i, I64x3, I64x3, I64x3, P(i)]),
        "stencil_subdomain_quantity": (i, [vp, i, i, I64x3, P(vp), P(vp), P(sz)]),
        "stencil_subdomain_to_host": (i, [vp, i, i, I64x3, I64x3, vp]),
        "stencil_subdomain_from_host": (i, [vp, i, i, I64x3, I64x3, vp]),
        "stencil_subdomain_interior": (i, [vp, i, I64x3, I64x3]),
        "stencil_subdomain_exterior": (i, [vp, i, (i64 * 3) * 6, (i64 * 3) * 6]),
        "stencil_jacobi_create": (vp, [u64, u64, u64, i, i, i, P(i), i, i]),
        "stencil_jacobi_destroy": (None, [vp]),
        "stencil_jacobi_init": (i, [vp]),
        "stencil_jacobi_init_random": (i, [vp, C.c_uint]),
        "stencil_jacobi_random_value": (C.c_float, [C.c_uint, i64, i64, i64, i64, i64, i64]),
        "stencil_jacobi_fingerprint": (u64, [vp]),
        "stencil_jacobi_step": (i, [vp]),
        "stencil_jacobi_sync": (i, [vp]),
        "stencil_jacobi_run": (dbl, [vp, i]),
        "stencil_jacobi_checksum": (dbl, [vp]),
        "stencil_jacobi_probe": (i, [vp, i64, i64, i64, P(C.c_float)]),
        "stencil_jacobi_launches": (u64, [vp]),
        "stencil_jacobi_mode": (i, [vp]),
        "stencil_jacobi_domain": (vp, [vp]),
        "stencil_jacobi_step_io": (i, [vp, vp, sz, vp]),
        "stencil_jacobi_host_step": (i, [vp, vp, i64, i64, i64]),
        "stencil_astaroth_create": (vp, [u64, u64, u64, i, i, i, i, i, i, P(i), i, dbl]),
        "stencil_astaroth_destroy": (None, [vp]),
        "stencil_astaroth_init": (i, [vp]),
        "stencil_astaroth_step": (i, [vp]),
        "stencil_astaroth_sync": (i, [vp]),
        "stencil_astaroth_run": (dbl, [vp, i, P(dbl)]),
        "stencil_astaroth_run_timed": (dbl, [vp, i, P(dbl), P(dbl)]),
        "stencil_astaroth_launches": (u64, [vp]),
        "stencil_astaroth_gather": (i, [vp, i, vp, u64]),
        "stencil_astaroth_host_step": (i, [vp, vp, i64, i64, i64, i, i, dbl]),
        "stencil_astaroth_domain": (vp, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(l, name)
        fn.restype = res
        fn.argtypes = args
    return l


def lib():
    """The loaded shared library (loaded once). Raises LibraryNotBuilt when the in-tree build is missing: there is
    deliberately no pure-Python fallback — on a GPU box a missing extension must fail loudly."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(_LIB_PATH):
                raise LibraryNotBuilt(
                    f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(or cmake -S . -B build -G Ninja && ninja -C build)")
            _lib = _declare(C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL))
    return _lib


def i3(v):
    return I64x3(int(v[0]), int(v[1]), int(v[2]))


def radius27(r):
    """Accepts an int (constant radius), a dict {(x,y,z): r} (missing = 0) or a 27-sequence."""
    out = U64x27()
    if isinstance(r, int):
        for k in range(27):
            out[k] = r
    elif isinstance(r, dict):
        for (x, y, z), v in r.items():
            out[((z + 1) * 3 + (y + 1)) * 3 + (x + 1)] = int(v)
    else:
        for k in range(27):
            out[k] = int(r[k])
    return out


def face_radius(r):
    return {d: r for d in [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]}
