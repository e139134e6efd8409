"""Runs the native GPU test binary (test/test_cuda_*.cu) on a B200: kernels, packers, LocalDomain, full exchanges
with every transport (2..8 subdomains oversubscribed on GPU 0 and, when present, all GPUs), the Jacobi engine."""
import os

import pytest

from conftest import run_native

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("group", ["[accessor]", "[align]", "[allocator]", "[array]", "[topo]", "[rcstream]", "[pack]", "[translate]",
                                   "[packer]", "[tx]", "[local_domain]", "[exchange]", "[jacobi]", "[astaroth]", "[derivative]"])
def test_native_cuda_group(bin_dir, group, tmp_path):
    r = run_native(os.path.join(bin_dir, "test_cuda"), group, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-6000:]
    assert " 0 failed" in r.stdout and "\n0 test cases" not in "\n" + r.stdout
