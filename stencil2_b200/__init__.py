"""stencil2_b200 — Python face of libstencil_b200.so, the Blackwell-native 3-D stencil halo-exchange library.

Everything heavy is native (C++17 + hand-written sm_100a CUDA in ``src/``); this package only loads the shared
library through its C ABI (``include/stencil_c.h``) and wraps it in a few small classes:

* :mod:`stencil2_b200.parallel`  partition / QAP placement / process group (CPU only, no GPU needed)
* :mod:`stencil2_b200.ops`       standalone pack / unpack / translate / reductions on device pointers
* :mod:`stencil2_b200.models`    DistributedDomain and the Jacobi3D solver
* :mod:`stencil2_b200.utils`     statistics, clock sampling, build helper

The library must have been built in-tree first (``python -c "import __graft_entry__ as g; g.build()"`` or CMake).
"""
from ._lib import lib, library_path, LibraryNotBuilt  # noqa: F401

__all__ = ["lib", "library_path", "LibraryNotBuilt", "version"]


def version() -> str:
    return lib().stencil_version().decode()
