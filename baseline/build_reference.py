#!/usr/bin/env python
"""Build the UNMODIFIED reference (mengshanfeng/stencil-2) into baseline/_ref for the `--impl reference` arm.

The prescribed `pip install ... /root/reference` cannot work: the reference is a CMake C++/CUDA project with no
setup.py / pyproject.toml (pip: "Directory '/root/reference' is not installable"). So its sources are copied verbatim
to baseline/_ref/src and compiled with nvcc for sm_100 with the reference's own Release flags. Two things the
image lacks are supplied from outside the source tree, without touching a reference file:
  * MPI      -> baseline/mpi_shim/mpi.h (single rank; see the header for what that can and cannot run)
  * NVTX v1  -> CUDA 12.9 ships only the header-only nvtx3; -I<cuda>/include/nvtx3 resolves <nvToolsExt.h>
libnvidia-ml is linked from the toolkit's stub (the real one comes with the driver on the GPU box).
/root/reference exists only on the build machine, so this runs there and baseline/_ref travels to the GPU box."""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")

APPS = {  # target -> (sources relative to the reference root, extra defines)
    "jacobi3d": (["bin/jacobi3d.cu", "bin/statistics.cpp"], []),
    "bench-pack": (["bin/bench_pack.cu"], []),
    "bench-exchange": (["bin/bench_exchange.cu", "bin/statistics.cpp"], []),
    "astaroth-sim": (["bin/astaroth_sim.cu"], []),
    "weak": (["bin/weak.cu"], ["-DSTENCIL_EXCHANGE_STATS=1"]),
    "strong": (["bin/strong.cu"], ["-DSTENCIL_EXCHANGE_STATS=1"]),
    "weak-exchange": (["bin/weak_exchange.cu"], ["-DSTENCIL_EXCHANGE_STATS=1"]),
    "bench-qap": (["bin/bench_qap.cu"], []),  # CPU only: placement solver timings (BASELINE config 1)
}
LIB = ["src/gpu_topology.cpp", "src/local_domain.cu", "src/rcstream.cpp", "src/stencil.cu"]


def sh(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout[-4000:])
        raise RuntimeError("reference build failed")
    return r.stdout


def build(force=False):
    src = os.path.join(OUT, "src")
    stamp = os.path.join(OUT, "BUILD_OK")
    if os.path.exists(stamp) and not force:
        return OUT
    if not os.path.isdir(REF):
        raise RuntimeError(f"{REF} not present on this machine and {OUT} has no finished build")
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    os.makedirs(OUT)
    shutil.copytree(REF, src, ignore=shutil.ignore_patterns(".git"))
    base = [os.path.join(CUDA, "bin", "nvcc"), "-std=c++14", "-O3", "-DNDEBUG", "-lineinfo", "-gencode", "arch=compute_100,code=sm_100",
            "--expt-extended-lambda", "-Wno-deprecated-gpu-targets", "-w", "-x", "cu",
            "-I" + os.path.join(src, "include"), "-I" + os.path.join(src, "thirdparty"), "-I" + os.path.join(HERE, "mpi_shim"),
            "-I" + os.path.join(CUDA, "include", "nvtx3"),
            "-DSTENCIL_USE_MPI=1", "-DSTENCIL_USE_CUDA=1", "-DSTENCIL_USE_CUDA_AWARE_MPI=0", "-DSTENCIL_SETUP_STATS=1", "-DSTENCIL_OUTPUT_LEVEL=2",
            "-DCATCH_CONFIG_NO_POSIX_SIGNALS"]
    objdir = os.path.join(OUT, "obj")
    bindir = os.path.join(OUT, "bin")
    os.makedirs(objdir)
    os.makedirs(bindir)
    jobs = []
    for variant, extra in (("default", []), ("stats", ["-DSTENCIL_EXCHANGE_STATS=1"])):
        for s in LIB:
            o = os.path.join(objdir, f"{variant}_{os.path.basename(s)}.o")
            jobs.append(base + extra + ["-c", os.path.join(src, s), "-o", o])
    for tgt, (srcs, extra) in APPS.items():
        for s in srcs:
            o = os.path.join(objdir, f"{tgt}_{os.path.basename(s)}.o")
            jobs.append(base + extra + ["-c", os.path.join(src, s), "-o", o])
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        list(ex.map(sh, jobs))
    for tgt, (srcs, extra) in APPS.items():
        variant = "stats" if extra else "default"
        objs = [os.path.join(objdir, f"{variant}_{os.path.basename(s)}.o") for s in LIB]
        objs += [os.path.join(objdir, f"{tgt}_{os.path.basename(s)}.o") for s in srcs]
        sh([os.path.join(CUDA, "bin", "nvcc"), "-gencode", "arch=compute_100,code=sm_100", "-o", os.path.join(bindir, tgt)] + objs +
           ["-L" + os.path.join(CUDA, "lib64", "stubs"), "-lnvidia-ml", "-lcudart"])
    shutil.rmtree(objdir)
    with open(stamp, "w") as f:
        f.write("reference built from %s with nvcc sm_100, mpi shim, nvtx3 headers\n" % REF)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
