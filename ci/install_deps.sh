#!/bin/bash
# The reference's ci/install_deps.sh downloaded CMake, OpenMPI and CUDA. This project builds offline against what the
# image has, so the step only verifies the toolchain and says what is missing.
set -euo pipefail
source "$(dirname "$0")/env.sh"
fail=0
need() { if ! command -v "$1" >/dev/null 2>&1; then echo "missing: $1 ($2)"; fail=1; fi; }
need nvcc "CUDA toolkit >= 12.8 with sm_100a support"
need cmake "CMake >= 3.24"
need ninja "Ninja"
need python "Python 3 with numpy and pytest for the binding tests"
if command -v nvcc >/dev/null 2>&1; then
  nvcc --list-gpu-arch 2>/dev/null | grep -q compute_100 || { echo "nvcc has no compute_100 target"; fail=1; }
fi
if command -v cmake >/dev/null 2>&1; then
  v=$(cmake --version | head -1 | awk '{print $3}')
  [ "$(printf '%s\n3.24\n' "$v" | sort -V | head -1)" = "3.24" ] || { echo "cmake $v is older than 3.24"; fail=1; }
fi
command -v mpicxx >/dev/null 2>&1 && echo "MPI found: configure with -DUSE_MPI=ON for multi-node runs" || echo "no MPI: process group = TCP (torchrun) or single process"
exit $fail
