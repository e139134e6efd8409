"""CPU-side checks of the Python binding: partition / QAP / placement through the C ABI, Statistics, the process
group over loopback with 2 real processes, and the host Jacobi loop."""
import os
import subprocess
import sys
import textwrap

import numpy as np

FACES1 = {(1, 0, 0): 1, (-1, 0, 0): 1, (0, 1, 0): 1, (0, -1, 0): 1, (0, 0, 1): 1, (0, 0, -1): 1}


def test_version_and_loader(built):
    import stencil2_b200
    assert stencil2_b200.version().startswith("stencil-b200")
    assert os.path.exists(stencil2_b200.library_path())


def test_rank_partition_matches_reference_cases(built):
    from stencil2_b200.parallel import RankPartition, set_partition_ties
    prev = set_partition_ties(True)  # the reference's cases: x is treated like the other axes
    try:
        _rank_partition_reference_cases(RankPartition)
    finally:
        set_partition_ties(prev)


def _rank_partition_reference_cases(RankPartition):
    p = RankPartition((10, 3, 1), 4)
    assert p.dim == (4, 1, 1)
    assert [p.subdomain((i, 0, 0)) for i in range(4)] == [((3, 3, 1), (0, 0, 0)), ((3, 3, 1), (3, 0, 0)), ((2, 3, 1), (6, 0, 0)), ((2, 3, 1), (8, 0, 0))]
    q = RankPartition((10, 14, 2), 9)
    assert q.dim == (3, 3, 1)
    assert q.subdomain((1, 1, 0))[1] == (4, 5, 0)
    assert q.subdomain((2, 2, 0))[1] == (7, 10, 0)


def test_node_partition_baseline_configs(built):
    from stencil2_b200.parallel import NodePartition, set_partition_ties
    # default: ties cut z, then y, then x (the contiguous axis last), and an x cut counts three times its size
    assert NodePartition((1024,) * 3, FACES1, 1, 8).dim == (1, 2, 4)
    p2 = NodePartition((645,) * 3, FACES1, 1, 2)
    assert p2.dim == (1, 1, 2) and p2.subdomain((0, 0, 0))[0] == (645, 645, 323) and p2.subdomain((0, 0, 1))[0] == (645, 645, 322)
    p4 = NodePartition((813,) * 3, FACES1, 1, 4)
    assert p4.dim == (1, 2, 2) and p4.subdomain((0, 1, 1))[0] == (813, 406, 406)
    # the reference's order on request
    prev = set_partition_ties(True)
    try:
        assert prev is False
        assert NodePartition((1024,) * 3, FACES1, 1, 8).dim == (2, 2, 2)
        p2 = NodePartition((645,) * 3, FACES1, 1, 2)
        assert p2.subdomain((0, 0, 0))[0] == (323, 645, 645) and p2.subdomain((1, 0, 0))[0] == (322, 645, 645)
        p4 = NodePartition((813,) * 3, FACES1, 1, 4)
        assert p4.dim == (2, 2, 1) and p4.subdomain((1, 1, 0))[0] == (406, 406, 813)
        uneven = {d: 1 for d in [(x, y, z) for x in (-1, 0, 1) for y in (-1, 0, 1) for z in (-1, 0, 1)]}
        uneven[(1, 0, 0)] = 2
        assert NodePartition((1024,) * 3, uneven, 1, 2).dim == (1, 2, 1)
        two = NodePartition((1024,) * 3, 3, 2, 4)
        assert two.sys_dim == (2, 1, 1) and two.node_dim == (1, 2, 2)
    finally:
        set_partition_ties(prev)


def test_qap_pinned_permutations(built):
    from stencil2_b200.parallel import qap_solve
    inf = float("inf")
    bw = [[900, 75, 64, 64], [75, 900, 64, 64], [64, 64, 900, 75], [64, 64, 75, 900]]
    comm = [[7, 5, 10, 1], [5, 7, 1, 10], [10, 1, 7, 5], [1, 10, 5, 7]]
    dist = [[(inf if v == 0 else 1.0 / v) for v in row] for row in bw]
    f, c = qap_solve(comm, dist, exact=True)
    assert f == [0, 2, 1, 3]
    g, c2 = qap_solve(comm, dist, exact=False)
    assert g == [3, 1, 2, 0]
    assert abs(c - c2) < 1e-12


def test_placement_512_cubed_on_synthetic_8gpu_topologies(built):
    """BASELINE config 1: 512^3 partition + QAP placement on synthetic 8-GPU NVLink topologies, no GPU."""
    from stencil2_b200.parallel import place_node_aware, set_partition_ties
    prev = set_partition_ties(True)  # the counts below are for the reference's 2x2x2 grid
    try:
        _placement_512_cases(place_node_aware)
    finally:
        set_partition_ties(prev)
    # default tie order: 512x512x1024 on 8 GPUs becomes 1x2x4 (x rows stay whole); still a valid assignment
    isl = [[100.0 if i == j else (25.0 if i % 2 == j % 2 else 1.0) for j in range(8)] for i in range(8)]
    cuda, dim = place_node_aware((512, 512, 1024), 3, isl)
    assert dim == (1, 2, 4) and sorted(cuda) == list(range(8))


def _placement_512_cases(place_node_aware):
    n = 8
    uniform = [[10.0 if i == j else 1.0 for j in range(n)] for i in range(n)]
    cuda, dim = place_node_aware((512, 512, 512), 3, uniform)
    assert dim == (2, 2, 2) and cuda == list(range(8))  # NVSwitch: identity
    # two islands of four GPUs with interleaved numbering: {0,2,4,6} and {1,3,5,7}
    isl = [[100.0 if i == j else (25.0 if i % 2 == j % 2 else 1.0) for j in range(n)] for i in range(n)]
    cuda, dim = place_node_aware((512, 512, 1024), 3, isl)
    # the heaviest interfaces must stay inside an island: subdomains that are x- or y-neighbours (512x1024 / 2 faces)
    sub = lambda x, y, z: cuda[x + dim[0] * (y + dim[1] * z)]
    assert sorted(cuda) == list(range(8))
    heavy_cross = 0
    for z in range(dim[2]):
        for y in range(dim[1]):
            for x in range(dim[0]):
                for dx, dy, dz in ((1, 0, 0), (0, 1, 0), (0, 0, 1)):
                    a = sub(x, y, z)
                    b = sub((x + dx) % dim[0], (y + dy) % dim[1], (z + dz) % dim[2])
                    if a != b and a % 2 != b % 2:
                        heavy_cross += 1
    # a 2x2x2 grid cut into two islands of four always has exactly one axis crossing: 4 pairs x 1 axis
    # (each counted from both ends here) — any more means a heavier face than necessary went over the slow link
    assert heavy_cross == 8


def test_statistics_trimean(built):
    from stencil2_b200.utils import Statistics
    s = Statistics()
    for v in [4, 1, 3, 2, 5, 8, 7, 6]:
        s.insert(v)
    assert s.trimean() == (3 + 2 * 5 + 7) / 4 and s.med() == 4.5 and s.min() == 1 and s.max() == 8


def test_host_jacobi_step_properties(built):
    from stencil2_b200.models import jacobi_host_step
    a = np.full((20, 30, 60), 0.5, dtype=np.float32)
    b = jacobi_host_step(a)
    assert b.shape == a.shape
    assert b[10, 15, 20] == 1.0 and b[10, 15, 40] == 0.0  # hot / cold centres (z=Z/2, y=Y/2, x=X/3 | 2X/3)
    assert b[0, 0, 0] == 0.5
    # radius X/10 = 6, truncated distance: a cell at distance 6.9 is inside, at 7.0 is outside
    assert b[10, 15, 26] == 1.0 and b[10, 15, 27] != 1.0


def test_host_astaroth_step_against_numpy(built):
    """The host reference of the 19-point radius-3 step is the 6th-order Laplacian: check it against numpy.roll."""
    from stencil2_b200.models import astaroth_host_step
    rng = np.random.default_rng(3)
    a = rng.standard_normal((9, 10, 12))
    dt = 0.01
    b = astaroth_host_step(a, stencil=1, fp64=True, dt=dt)
    lap = np.zeros_like(a)
    for axis in range(3):
        lap += 3 * (-49.0 / 18.0) * a / 3
        for k, c in ((1, 1.5), (2, -3.0 / 20.0), (3, 1.0 / 90.0)):
            lap += c * (np.roll(a, k, axis) + np.roll(a, -k, axis))
    assert np.allclose(b, a + dt * lap, rtol=0, atol=1e-12)
    # the reference's stand-in: plain 7-point average of the six face neighbours
    c = astaroth_host_step(a, stencil=0, fp64=True)
    avg = sum(np.roll(a, s, ax) for ax in range(3) for s in (1, -1)) / 6
    assert np.allclose(c, avg, rtol=0, atol=1e-13)
    # cross-derivative variant: 6th-order mixed second derivatives along the diagonals of the three planes plus the
    # d3/dxdydz corner term, weight 0.25 (array axes are [z, y, x]; f(x+k) is np.roll(a, -k, axis=2))
    d = astaroth_host_step(a, stencil=2, fp64=True, dt=dt)
    sh = lambda dx, dy, dz: np.roll(a, (-dz, -dy, -dx), axis=(0, 1, 2))
    cross = np.zeros_like(a)
    for k, m in ((1, 270.0 / 720.0), (2, -27.0 / 720.0), (3, 2.0 / 720.0)):
        cross += m * (sh(k, k, 0) - sh(k, -k, 0) - sh(-k, k, 0) + sh(-k, -k, 0))
        cross += m * (sh(k, 0, k) - sh(-k, 0, k) - sh(k, 0, -k) + sh(-k, 0, -k))
        cross += m * (sh(0, k, k) - sh(0, -k, k) - sh(0, k, -k) + sh(0, -k, -k))
    cross += (270.0 / 720.0) * sum(sx * sy * sz * sh(sx, sy, sz) for sx in (1, -1) for sy in (1, -1) for sz in (1, -1))
    assert np.allclose(d, a + dt * (lap + 0.25 * cross), rtol=0, atol=1e-12)
    assert np.abs(cross).max() > 0.1  # the terms are really there


def test_two_process_tcp_group(built, tmp_path):
    """world_size 2 on this machine: the native Comm joins through MASTER_ADDR/MASTER_PORT like under torchrun."""
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent(f"""
        import sys
        sys.path.insert(0, {built!r})
        from stencil2_b200.parallel import comm
        r, n = comm.init_from_env()
        assert n == 2 and comm.rank() == r
        assert comm.allreduce_max(float(r)) == 1.0
        comm.barrier()
        comm.finalize()
        print("ok", r)
    """))
    port = 24000 + os.getpid() % 2000
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOCAL_RANK=str(r))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "ok 0" in outs[0] and "ok 1" in outs[1]


def test_c_abi_header_library_and_binding_agree(built):
    """Every function declared in include/stencil_c.h is exported by the library, and everything the ctypes layer
    declares exists in the header (so the three cannot drift apart silently)."""
    import re
    import subprocess
    from stencil2_b200 import library_path
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "stencil_c.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(stencil_[a-z0-9_]+)\s*\(", header))
    nm = subprocess.run(["nm", "-D", "--defined-only", library_path()], stdout=subprocess.PIPE, text=True, check=True).stdout
    exported = {l.split()[-1] for l in nm.splitlines() if " T " in l and "stencil_" in l}
    missing = sorted(d for d in declared if d not in exported)
    assert not missing, f"declared in stencil_c.h but not exported: {missing}"
    binding = open(os.path.join(root, "stencil2_b200", "_lib.py")).read()
    bound = set(re.findall(r'"(stencil_[a-z0-9_]+)"\s*:', binding))
    unknown = sorted(b for b in bound if b not in declared)
    assert not unknown, f"bound in _lib.py but not in stencil_c.h: {unknown}"
    assert len(bound) > 40


def test_bench_line_contract_and_weak_rule():
    """The JSON line bench.py printed on the GPU box (kept under profiles/) has every key the driver reads, and the
    weak-scaling rule gives the reference's domain sizes (645^3, 813^3, 1024^3)."""
    import importlib.util
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert [bench.weak_size(512, n) for n in (1, 2, 4, 8)] == [512, 645, 813, 1024]
    assert [bench.weak_size(256, n) for n in (2, 4, 8)] == [323, 406, 512]
    cfg = bench.shared_config(512, 8)
    assert cfg["global_batch"] == "1024x1024x1024 cells" and cfg["parallelism"] == "dd8"
    line_file = os.path.join(root, "profiles", "runs", "r2", "bench_n1.json")
    if not os.path.exists(line_file):
        return
    line = json.loads(open(line_file).read().strip().splitlines()[-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "clocks", "e2e", "gpu_launches"):
        assert key in line, key
    assert line["metric"] == bench.METRIC and line["higher_is_better"] is True and line["scaling"] == "weak"
    assert line["warmup"] >= 3 and line["gpu_launches"] >= line["steps"]
    for key in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert key in line["e2e"], key
    assert line["config"] == bench.shared_config(512, 1)
    assert line["verified"] is True
    # cells per second from the time per step, as printed
    assert abs(line["value"] - 512.0 ** 3 / (line["ms_per_step"] * 1e-3) / 1e9) < 0.01 * line["value"]
