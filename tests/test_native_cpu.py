"""Runs the native CPU test binary (test/test_cpu_*.cpp): value types, Mat2D, partition, QAP, placement on synthetic
topologies, tags, Statistics, argparse, the Comm layer over loopback TCP."""
import os

from conftest import run_native


def test_native_cpu_suite(bin_dir):
    r = run_native(os.path.join(bin_dir, "test_cpu"))
    assert r.returncode == 0, r.stdout[-4000:]
    assert " 0 failed" in r.stdout


def test_bench_qap_runs_on_cpu(bin_dir):
    exe = os.path.join(bin_dir, "bench-qap")
    if not os.path.exists(exe):
        import pytest
        pytest.skip("bench-qap not built yet")
    r = run_native(exe, "--max-n", "10", "--reps", "2")
    assert r.returncode == 0, r.stdout[-2000:]
    assert "size" in r.stdout


def test_pingpong_between_two_ranks(bin_dir):
    """pingpong (reference: bin/pingpong.cu over MPI_Send/Recv) over the TCP process group, two ranks on this host:
    one line per node pair with one time per message size 2^min .. 2^max."""
    import subprocess
    exe = os.path.join(bin_dir, "pingpong")
    port = 24100 + os.getpid() % 800
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   STENCIL_LOG_LEVEL="1")
        procs.append(subprocess.Popen([exe, "1", "--min", "0", "--max", "10", "--iters", "3", "--all-ranks"], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True, cwd="/tmp"))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert [p.returncode for p in procs] == [0, 0], outs
    fields = outs[0].split()
    assert len(fields) == 1 + 11 and all(float(v) > 0 for v in fields[1:]), outs[0]
