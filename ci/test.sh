#!/bin/bash
# CI test step (reference: ci/test.sh, which Travis never ran). CPU tier always; GPU tier when a device is present.
set -euo pipefail
cd "$(dirname "$0")/.."
./build/bin/test_cpu
python -m pytest tests -q -m "not gpu"
if ./build/bin/test_cuda "[align]" >/dev/null 2>&1; then
  python -m pytest tests -q -m gpu
  # sanitizer gate (the reference's CTest command is `cuda-memcheck test_cuda`): memcheck + racecheck over the tests of
  # the hand-written synchronisation; STENCIL_CI_SANITIZE=0 skips it on pools where compute-sanitizer is closed
  if [ "${STENCIL_CI_SANITIZE:-1}" = 1 ]; then
    tools/gpu_sanitize.sh
    if grep -q "rc=86" gpurun_out/sanitizer/summary.txt; then
      echo "compute-sanitizer is closed on this pool: gate skipped"
    elif grep -q "rc=[1-9]" gpurun_out/sanitizer/summary.txt; then
      echo "sanitizer gate failed"; exit 1
    fi
  fi
else
  echo "no CUDA device: GPU tier skipped"
fi
