"""Multi-process exchanges on a GPU box (the reference's `mpiexec -n 2 test_cuda_mpi` tier): N real processes joined
through the TCP process group, each owning subdomains, exchanging through every transport that is eligible between
processes. On a 1-GPU box the ranks share GPU 0 — the staged and colocated (CUDA IPC) transports carry the traffic;
with >= 2 GPUs each rank takes its own GPU and the direct NVLink path (IPC-mapped slabs + device flags) is used."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

STAGED, COLO, PEER, KERNEL, DIRECT = 1, 4, 8, 16, 32


def _launch(bin_dir, nranks, methods, extra=(), port_base=26000, env_extra=None):
    exe = os.path.join(bin_dir, "test_mp_exchange")
    port = port_base + os.getpid() % 1500
    procs = []
    for r in range(nranks):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(nranks), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   STENCIL_LOG_LEVEL="1")
        if env_extra:
            env.update(env_extra)
        procs.append(subprocess.Popen([exe, str(methods), *map(str, extra)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            p.kill()
            outs.append("TIMEOUT")
    return [p.returncode for p in procs], outs


@pytest.mark.parametrize("methods", [STAGED, STAGED | KERNEL, COLO | KERNEL | STAGED, STAGED | COLO | PEER | KERNEL | DIRECT])
def test_two_ranks(bin_dir, methods):
    rcs, outs = _launch(bin_dir, 2, methods, port_base=26000 + methods * 7)
    assert rcs == [0, 0], outs
    assert all(" ok " in o for o in outs), outs


def test_three_ranks_uneven_sizes(bin_dir):
    rcs, outs = _launch(bin_dir, 3, STAGED | COLO | KERNEL | DIRECT, extra=(31, 17, 13, 3, 2), port_base=27800)
    assert rcs == [0, 0, 0], outs


def test_two_ranks_posing_as_two_nodes(bin_dir):
    """STENCIL_FAKE_HOSTNAME makes each rank its own node: placement partitions across nodes and every message between
    the ranks has to take the staged (cross-node fallback) transport even though all methods are enabled."""
    exe_env = None
    port = 28900 + os.getpid() % 500
    exe = os.path.join(bin_dir, "test_mp_exchange")
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   STENCIL_LOG_LEVEL="1", STENCIL_FAKE_HOSTNAME=f"node{r}")
        procs.append(subprocess.Popen([exe, str(STAGED | COLO | PEER | KERNEL | DIRECT)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert [p.returncode for p in procs] == [0, 0], outs
    del exe_env


def _launch_app(bin_dir, app, nranks, args, port_base):
    exe = os.path.join(bin_dir, app)
    port = port_base + os.getpid() % 700
    procs = []
    for r in range(nranks):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(nranks), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   STENCIL_LOG_LEVEL="1")
        procs.append(subprocess.Popen([exe, *args], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd="/tmp"))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            p.kill()
            outs.append("TIMEOUT")
    return [p.returncode for p in procs], outs


@pytest.mark.parametrize("nranks", [2, 3])
def test_astaroth_engine_across_processes(bin_dir, nranks):
    """AstarothSim with one process per rank, checked against a host loop on every rank (astaroth-sim --check). With one
    GPU the ranks share it (staged / colocated transports, blocking schedule); with a GPU per rank the free-running
    loop with kernel-written x halos and flag-synchronised y / z pushes over IPC-mapped slabs is what runs."""
    args = ["--x", "64", "--y", "48", "--z", "66", "--fields", "3", "--double", "--star", "--iters", "8", "--check"]
    rcs, outs = _launch_app(bin_dir, "astaroth-sim", nranks, args, 30100 + 900 * nranks)
    assert rcs == [0] * nranks, outs
    assert "astaroth-check" in outs[0] and ",ok" in outs[0], outs
