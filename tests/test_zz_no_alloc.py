"""The allocation-free field sum (stencil_reduce_sum_f32_noalloc), called through the C ABI.

Some callers of this library cannot allocate device memory at the moment they want a field sum — a residual monitor
inside a fused, free-running iteration loop (cudaMalloc/cudaFree would synchronise the device under kernels that wait
on their neighbours), or a solver that has handed all of HBM to its subdomains. For them the library carries the
reduction's scratch in a fixed device array. This test checks that entry point against the allocating one
(stencil_reduce_sum_f32) on the same data, twice in a row, because the second call finds the first call's leftovers
in the shared scratch. It lives in a file that sorts last so that a `-x` run has already run everything else; it
must pass both on its own in a fresh process and after the other tests."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-9  # same tolerance as the allocating reduction's test in test_python_gpu.py


def _sum_alloc(L, i3, t, x0, ext, pitch):
    out = C.c_double()
    rc = L.stencil_reduce_sum_f32(t.data_ptr() + 4 * x0, i3(pitch), i3(ext), None, C.byref(out))
    assert rc == 0
    return out.value


def test_noalloc_sum_matches_allocating_sum_twice(built):
    import torch
    from stencil2_b200._lib import lib, i3
    L = lib()
    torch.manual_seed(5)
    # buffers are created by the TEST (torch) before the calls under test; the entry point itself allocates nothing
    a = torch.rand(33, 47, 200, dtype=torch.float32, device="cuda")
    b = torch.rand(17, 64, 512, dtype=torch.float32, device="cuda") * 3.0 - 1.0
    out = torch.full((2,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    cases = [(a, 3, (150, 47, 33), (200, 47, 33)), (b, 0, (512, 64, 17), (512, 64, 17))]
    for k, (t, x0, ext, pitch) in enumerate(cases):
        want = _sum_alloc(L, i3, t, x0, ext, pitch)
        rc = L.stencil_reduce_sum_f32_noalloc(t.data_ptr() + 4 * x0, i3(pitch), i3(ext), None, out.data_ptr() + 8 * k)
        assert rc == 0
        torch.cuda.synchronize()
        got = out[k].item()
        assert abs(got - want) <= REL_TOL * abs(want), (k, got, want)
        ref = t[:, :, x0:x0 + ext[0]].to(torch.float64).sum().item()
        assert abs(got - ref) <= 1e-9 * abs(ref) + 1e-6
    # and once more on the first input: same answer again, bit for bit (fixed summation order)
    again = torch.zeros(1, dtype=torch.float64, device="cuda")
    t, x0, ext, pitch = cases[0]
    assert L.stencil_reduce_sum_f32_noalloc(t.data_ptr() + 4 * x0, i3(pitch), i3(ext), None, again.data_ptr()) == 0
    torch.cuda.synchronize()
    assert again[0].item() == out[0].item()


def test_noalloc_rejects_what_it_cannot_serve(built, capfd):
    import torch
    from stencil2_b200._lib import lib, i3
    L = lib()
    t = torch.zeros(8, dtype=torch.float32, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    too_many_rows = (4, 1 << 20, 1 << 12)  # 2^32 rows > STENCIL_NOALLOC_MAX_ROWS
    assert L.stencil_reduce_sum_f32_noalloc(t.data_ptr(), i3(too_many_rows), i3(too_many_rows), None, out.data_ptr()) == -3
    assert "exceeds the supported maximum" in capfd.readouterr().err
    assert L.stencil_reduce_sum_f32_noalloc(None, i3((8, 1, 1)), i3((8, 1, 1)), None, out.data_ptr()) == -1
    assert L.stencil_reduce_sum_f32_noalloc(t.data_ptr(), i3((4, 1, 1)), i3((8, 1, 1)), None, out.data_ptr()) == -2
    torch.cuda.synchronize()
    assert out[0].item() == 0.0  # nothing was launched
