"""In-tree build of libstencil_b200.so, the apps and the native tests: CMake + Ninja, sm_100a only.
nvcc cross-compiles without a GPU, so this works on CPU-only machines."""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(build_dir=None, verbose=False, targets=None):
    build_dir = build_dir or os.path.join(ROOT, "build")
    cmake = shutil.which("cmake")
    ninja = shutil.which("ninja")
    if not cmake or not ninja:
        raise RuntimeError("cmake and ninja are required (both ship in this image's virtualenv)")
    env = dict(os.environ)
    env.setdefault("CUDACXX", "/usr/local/cuda/bin/nvcc")
    cache = os.path.join(build_dir, "CMakeCache.txt")
    if os.path.exists(cache):
        # a cache configured from another checkout path (e.g. the GPU box's scratch copy) cannot be reused
        with open(cache) as f:
            home = [l for l in f if l.startswith("CMAKE_HOME_DIRECTORY")]
        if not home or os.path.realpath(home[0].split("=", 1)[1].strip()) != os.path.realpath(ROOT):
            shutil.rmtree(build_dir)
    os.makedirs(build_dir, exist_ok=True)
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call([cmake, "-S", ROOT, "-B", build_dir, "-G", "Ninja", "-DCMAKE_BUILD_TYPE=Release"], env=env, stdout=out)
    cmd = [ninja, "-C", build_dir] + (list(targets) if targets else [])
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-8000:])
        raise RuntimeError("native build failed")
    if verbose:
        sys.stdout.write(r.stdout[-2000:])
    return os.path.join(ROOT, "stencil2_b200", "lib", "libstencil_b200.so")


if __name__ == "__main__":
    print(build(verbose=True))
