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
