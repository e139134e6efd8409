import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def _ensure_built():
    lib = os.path.join(ROOT, "stencil2_b200", "lib", "libstencil_b200.so")
    bins = [os.path.join(ROOT, "build", "bin", b) for b in ("test_cpu", "test_cuda", "jacobi3d")]
    if os.path.exists(lib) and all(os.path.exists(b) for b in bins):
        return
    import __graft_entry__
    __graft_entry__.build()


@pytest.fixture(scope="session")
def built():
    _ensure_built()
    return ROOT


@pytest.fixture(scope="session")
def bin_dir(built):
    return os.path.join(built, "build", "bin")


def run_native(exe, *args, timeout=900, env=None, cwd=None):
    e = dict(os.environ)
    e.setdefault("STENCIL_LOG_LEVEL", "2")
    if env:
        e.update(env)
    return subprocess.run([exe, *args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, env=e, cwd=cwd)
