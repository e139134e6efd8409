"""The apps keep the reference's command lines and output formats. Tiny sizes: this checks formats and
that every binary runs, not speed."""
import os
import re

import pytest

from conftest import run_native

pytestmark = pytest.mark.gpu


def _run(bin_dir, tmp_path, name, *args):
    r = run_native(os.path.join(bin_dir, name), *args, cwd=str(tmp_path), env={"STENCIL_LOG_LEVEL": "1"}, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    return r.stdout


def test_jacobi3d_csv_line_and_paraview(bin_dir, tmp_path):
    out = _run(bin_dir, tmp_path, "jacobi3d", "48", "32", "24", "-n", "6", "--paraview", "--prefix", "pv_", "--period", "3")
    line = [l for l in out.splitlines() if l.startswith("jacobi3d,")][-1].split(",")
    assert len(line) == 9
    assert line[1] == "staged/colo/peer/kernel/direct" and int(line[2]) == 1 and int(line[3]) >= 1
    assert float(line[7]) > 0 and float(line[8]) >= float(line[7])
    files = sorted(os.listdir(tmp_path))
    assert "pv_jacobi3d_init_0.txt" in files and "pv_jacobi3d_final_0.txt" in files and "pv_jacobi3d_3_0.txt" in files
    assert "plan_0.txt" in files
    with open(tmp_path / "pv_jacobi3d_final_0.txt") as f:
        assert f.readline().strip() == "Z,Y,X,d"
        assert re.match(r"^\d+,\d+,\d+,-?\d+\.\d{6}$", f.readline().strip())
    # method selection flags and the reference schedules
    for extra in (["--staged"], ["--peer", "--kernel"], ["--no-overlap"], ["--no-fused", "--trivial"]):
        out = _run(bin_dir, tmp_path, "jacobi3d", "40", "24", "20", "-n", "3", *extra)
        assert any(l.startswith("jacobi3d,") for l in out.splitlines())


def test_weak_strong_lines(bin_dir, tmp_path):
    for app in ("weak", "strong"):
        out = _run(bin_dir, tmp_path, app, "40", "32", "24", "4", "--kernel", "--staged")
        f = [l for l in out.splitlines() if l.startswith("weak,")][-1].split(",")
        assert len(f) == 23 and f[1] == "staged/kernel"
        assert int(f[10]) == 4 and float(f[21]) > 0  # iters, exchange seconds
        assert int(f[6]) + int(f[7]) + int(f[8]) + int(f[9]) > 0  # bytes per exchange
    out = _run(bin_dir, tmp_path, "weak-exchange", "40", "32", "24", "4")
    f = [l for l in out.splitlines() if l.startswith("weak,")][-1].split(",")
    assert len(f) == 15 and f[1].endswith("/all") and float(f[14]) > 0
    assert any(l.startswith("weak-exchange-device,") for l in out.splitlines())


def test_bench_exchange_table(bin_dir, tmp_path):
    out = _run(bin_dir, tmp_path, "bench-exchange", "--x", "32", "--y", "24", "--z", "16", "--iters", "3", "--q", "2", "--fr", "2", "--er", "1")
    lines = out.strip().splitlines()
    assert lines[0] == "name,count,trimean (S),trimean (B/s),stddev,min,avg,max"
    names = [l.split(",")[0] for l in lines[1:]]
    assert names == ["32-24-16/px/2", "32-24-16/x/2", "32-24-16/faces/2", "32-24-16/face&edge/2/1", "32-24-16/uniform/2"]
    assert all(len(l.split(",")) == 8 and int(l.split(",")[1]) == 3 for l in lines[1:])


def test_bench_pack_lines(bin_dir, tmp_path):
    out = _run(bin_dir, tmp_path, "bench-pack", "--iters", "3", "--x", "64", "--y", "48", "--z", "32")
    lines = out.strip().splitlines()
    assert len(lines) == 3
    assert lines[0].split()[0] == "[3,48,32]" and lines[0].split()[1] == "[1,0,0]" and int(lines[0].split()[2]) == 3 * 48 * 32 * 4
    assert lines[2].split()[1] == "[0,0,1]"


def test_astaroth_sim(bin_dir, tmp_path):
    out = _run(bin_dir, tmp_path, "astaroth-sim", "--x", "40", "--y", "32", "--z", "24")  # the reference's workload
    f = [l for l in out.splitlines() if l.startswith("astaroth-sim,")][-1].split(",")
    assert f[7:11] == ["1", "fp32", "avg7", "5"]
    out = _run(bin_dir, tmp_path, "astaroth-sim", "--x", "40", "--y", "32", "--z", "24", "--fields", "8", "--double", "--star", "--iters", "3", "--peer", "--kernel")
    f = [l for l in out.splitlines() if l.startswith("astaroth-sim,")][-1].split(",")
    assert f[1] == "peer/kernel" and f[7:11] == ["8", "fp64", "star19", "3"] and float(f[13]) > 0


def test_astaroth_sim_cross_terms_check(bin_dir, tmp_path):
    """--cross reads edge and corner halos; --check compares with the host loop after steps in which every halo comes from
    the kernels' own pushes."""
    out = _run(bin_dir, tmp_path, "astaroth-sim", "--x", "70", "--y", "44", "--z", "50", "--fields", "2", "--double", "--cross", "--iters", "6", "--check")
    line = [l for l in out.splitlines() if l.startswith("astaroth-check,")][-1]
    assert line.endswith(",ok"), out


def test_jacobi3d_verify_mode(bin_dir, tmp_path):
    """jacobi3d --verify: fused kernel vs plain kernel + exchange() from a random field, fingerprints must agree; with one
    GPU held back before every iteration the flag protocol still has to keep the iterations apart."""
    out = _run(bin_dir, tmp_path, "jacobi3d", "200", "70", "40", "--exact-size", "--verify", "12", "--tile-z", "8", "--delay-us", "200")
    line = [l for l in out.splitlines() if l.startswith("jacobi3d-verify,")][-1].split(",")
    assert line[-1] == "ok" and line[-2] == line[-3] and line[6] == "12", out


def test_topology_benchmarks_run(bin_dir, tmp_path):
    out = _run(bin_dir, tmp_path, "bench-alltoallv", "--iters", "2", "--cap-mib", "4")
    assert out.splitlines()[0] == "stencil" and "direct-store" in out and "All-to-all 8MiB" in out
    out = _run(bin_dir, tmp_path, "measure-buf-exchange", "--rounds", "2", "--measures", "2")
    assert "final MiB" in out and "GB/s" in out


def test_example_heat_conserves_the_mean(bin_dir, tmp_path):
    """examples/heat.cu: a user-written stencil with interior / exchange_async / exterior overlap."""
    out = _run(bin_dir, tmp_path, "example-heat", "64", "48", "40", "12")
    line = [l for l in out.splitlines() if l.startswith("mean temperature")][-1]
    assert abs(float(line.split()[-1]) - 0.5) < 1e-3, out
