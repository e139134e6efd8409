"""Process group. Under ``torchrun`` (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT in the environment) the native
library joins a TCP group by itself on first use; ``init_from_env`` just makes that explicit and lets callers pick
the port. With a lone process everything is a no-op."""
import os

from .._lib import lib


def init_from_env(port=None):
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws <= 1:
        lib().stencil_comm_init_tcp(0, 1, b"127.0.0.1", 0)
        return 0, 1
    rk = int(os.environ["RANK"])
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    if port is None:
        port = int(os.environ.get("STENCIL_PORT", int(os.environ.get("MASTER_PORT", "29500")) + 17))
    lib().stencil_comm_init_tcp(rk, ws, addr.encode(), int(port))
    return rk, ws


def rank():
    return lib().stencil_comm_rank()


def size():
    return lib().stencil_comm_size()


def barrier():
    lib().stencil_comm_barrier()


def allreduce_max(v):
    return lib().stencil_comm_allreduce_max(float(v))


def finalize():
    lib().stencil_comm_finalize()
