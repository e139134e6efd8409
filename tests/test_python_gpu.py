"""GPU checks through the Python binding (the same path bench.py and smoke() use)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_smoke_entry(built):
    import __graft_entry__
    __graft_entry__.smoke()


def test_jacobi_modes_agree_with_host(built):
    from stencil2_b200.models import Jacobi3D, jacobi_host_step
    size = (70, 20, 18)
    ref = np.full((size[2], size[1], size[0]), 0.5, dtype=np.float32)
    for _ in range(5):
        ref = jacobi_host_step(ref)
    for mode in (Jacobi3D.FUSED, Jacobi3D.OVERLAP, Jacobi3D.NO_OVERLAP):
        j = Jacobi3D(*size, gpus=[0, 0, 0], mode=mode, tile_z=8)
        for _ in range(5):
            j.step()
        assert np.array_equal(j.gather(), ref), f"mode {mode}"
        assert abs(j.checksum() - float(ref.astype(np.float64).sum())) < 1e-6 * ref.size
        j.close()


def test_astaroth_engine_against_host(built):
    from stencil2_b200.models import AstarothSim, astaroth_host_step
    for size, fp64, tol in (((140, 24, 20), True, 1e-13), ((140, 24, 20), False, 5e-6)):
        for overlap in (None, True):
            sim = AstarothSim(*size, fields=2, fp64=fp64, stencil=AstarothSim.STAR19, overlap=overlap, gpus=[0, 0])
            sim.init()
            ref = sim.gather(1)
            assert np.isfinite(ref).all() and ref.std() > 0.1
            for _ in range(3):
                sim.step()
                ref = astaroth_host_step(ref, stencil=1, fp64=fp64, dt=sim.dt)
            sim.sync()
            assert np.abs(sim.gather(1) - ref).max() < tol
            assert sim.launches > 0
            sim.close()


def test_jacobi_random_field_fingerprint_and_cross_stencil(built):
    """What bench.py's verification relies on: a random initial field that is a function of the global coordinate, and a
    field fingerprint that does not depend on partition or schedule; and the cross-derivative Astaroth stencil (reads
    edge and corner halos) against the host reference, with every halo coming from the kernels' pushes after step 1."""
    from stencil2_b200._lib import lib
    from stencil2_b200.models import Jacobi3D, AstarothSim, astaroth_host_step, jacobi_host_step
    size = (133, 40, 30)
    ref = np.empty((size[2], size[1], size[0]), dtype=np.float32)
    L = lib()
    for z in range(0, size[2], 7):
        for y in range(0, size[1], 5):
            for x in range(0, size[0], 11):
                ref[z, y, x] = L.stencil_jacobi_random_value(7, x, y, z, *size)
    prints = []
    fields = []
    for mode, gpus in ((Jacobi3D.FUSED, [0]), (Jacobi3D.NO_OVERLAP, [0, 0]), (Jacobi3D.FUSED, [0, 0, 0, 0])):
        j = Jacobi3D(*size, gpus=gpus, mode=mode, tile_z=8)
        j.init_random(7)
        got = j.gather()
        assert all(got[z, y, x] == ref[z, y, x] for z in range(0, size[2], 7) for y in range(0, size[1], 5) for x in range(0, size[0], 11))
        assert 0.0 <= got.min() and got.max() < 1.0 and got.std() > 0.2
        for _ in range(6):
            j.step()
        prints.append(j.fingerprint())
        fields.append(j.gather())
        j.close()
    assert prints[0] == prints[1] == prints[2] and prints[0] != 0
    assert np.array_equal(fields[0], fields[1]) and np.array_equal(fields[0], fields[2])
    host = got
    for _ in range(6):
        host = jacobi_host_step(host)
    assert np.array_equal(host, fields[0])

    sim = AstarothSim(70, 44, 50, fields=2, fp64=True, stencil=AstarothSim.STAR19_CROSS, overlap=False, gpus=[0, 0, 0, 0])
    sim.init()
    ref = sim.gather(0)
    for _ in range(7):
        sim.step()
        ref = astaroth_host_step(ref, stencil=2, fp64=True, dt=sim.dt)
    sim.sync()
    assert np.abs(sim.gather(0) - ref).max() < 1e-12
    t, ex, dev = sim.run_timed(3)
    assert t > 0 and dev > 0 and dev <= t * 1.5
    sim.close()


def test_domain_grid_and_timed_exchange(built):
    from stencil2_b200.models import DistributedDomain, Method
    dd = DistributedDomain(40, 30, 20)
    dd.set_radius(2)
    dd.add_data(np.float32, "a")
    dd.add_data(np.float64, "b")
    dd.set_methods(Method.ALL)
    dd.set_gpus([0, 0, 0, 0])
    dd.set_plan_file(False)
    dd.realize()
    gx, gy, gz = dd.grid
    assert gx * gy * gz == 4 and gx == 1  # x is the expensive axis: cut last
    host_s, dev_s = dd.time_exchange(5)
    assert host_s > 0 and dev_s > 0 and dev_s <= host_s * 1.5
    dd.close()


def test_exchange_async_chained_to_a_torch_stream(built):
    """DistributedDomain.exchange_async(): halos of a periodic 2-subdomain field are filled while the host never waits
    for the exchange; a torch stream is chained to it through the raw CUDA events in both directions."""
    import torch
    from stencil2_b200.models import DistributedDomain, Method
    dd = DistributedDomain(16, 12, 10)
    dd.set_radius(1)
    q = dd.add_data(np.float32, "f")
    dd.set_methods(Method.ALL)
    dd.set_gpus([0, 0])
    dd.set_plan_file(False)
    dd.realize()
    assert dd.exchange_device_only
    g = np.arange(16 * 12 * 10, dtype=np.float32).reshape(10, 12, 16)
    subs = dd.subdomains
    for s in subs:
        (ox, oy, oz), (sx, sy, sz) = s.origin, s.size
        s.from_host(q, np.full((s.raw_size[2], s.raw_size[1], s.raw_size[0]), -1, np.float32))
        s.from_host(q, g[oz:oz + sz, oy:oy + sy, ox:ox + sx], pos=(1, 1, 1))
    stream = torch.cuda.Stream()
    mine = torch.cuda.Event()
    for rounds in range(2):
        mine.record(stream)
        for s in subs:
            dd.exchange_after(s.index, mine.cuda_event)
        dd.exchange_async()
        for s in subs:
            ev = dd.exchange_done_event(s.index)
            assert ev
            try:
                from cuda.bindings import runtime as cudart
            except ImportError:
                from cuda import cudart
            (err,) = cudart.cudaStreamWaitEvent(stream.cuda_stream, ev, 0)
            assert int(err) == 0
        stream.synchronize()
        for s in subs:
            (ox, oy, oz), (sx, sy, sz) = s.origin, s.size
            got = s.to_host(q, np.float32)
            zz, yy, xx = np.meshgrid(np.arange(-1, sz + 1), np.arange(-1, sy + 1), np.arange(-1, sx + 1), indexing="ij")
            want = g[(oz + zz) % 10, (oy + yy) % 12, (ox + xx) % 16]
            assert np.array_equal(got, want)
    dd.exchange_wait()
    dd.close()


def test_ops_pack_and_reduce_against_numpy(built):
    import torch
    from stencil2_b200.ops import pack, unpack, reduce_sum_f32
    src = torch.arange(40 * 12 * 9, dtype=torch.float32, device="cuda").reshape(9, 12, 40)
    dst = torch.zeros(4, 5, 16, dtype=torch.float32, device="cuda")
    pack(dst.data_ptr(), src.data_ptr(), (40, 12, 9), (8, 3, 2), (16, 5, 4), 4)
    torch.cuda.synchronize()
    assert torch.equal(dst, src[2:6, 3:8, 8:24])
    back = torch.zeros_like(src)
    unpack(back.data_ptr(), (40, 12, 9), (8, 3, 2), (16, 5, 4), dst.data_ptr(), 4)
    torch.cuda.synchronize()
    assert torch.equal(back[2:6, 3:8, 8:24], dst) and back.sum().item() == dst.sum().item()
    field = torch.rand(33, 47, 200, dtype=torch.float32, device="cuda")
    got = reduce_sum_f32(field.data_ptr() + 4 * 3, (200, 47, 33), (150, 47, 33))
    want = field[:, :, 3:153].to(torch.float64).sum().item()
    assert abs(got - want) <= 1e-9 * abs(want)


def test_bench_contract_small(built):
    """bench.py prints one JSON line with the contract's keys (tiny size so it finishes in seconds)."""
    r = subprocess.run([sys.executable, os.path.join(built, "bench.py"), "--gpus", "1", "--steps", "4", "--warmup", "3", "--size", "96"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "clocks", "e2e", "gpu_launches"):
        assert k in out, k
    assert out["value"] > 0 and out["e2e"]["value"] > 0 and out["gpu_launches"] >= 4
    assert out["e2e"]["h2d_bytes_per_step"] > 0 and out["e2e"]["d2h_bytes_per_step"] > 0
