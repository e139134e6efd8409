#!/bin/bash
# CI test step (reference: ci/test.sh, which Travis never ran). CPU tier always; GPU tier when a device is present.
set -euo pipefail
cd "$(dirname "$0")/.."
./build/bin/test_cpu
python -m pytest tests -q -m "not gpu"
if ./build/bin/test_cuda "[align]" >/dev/null 2>&1; then
  # memcheck is optional: some pools close compute-sanitizer (STENCIL_CI_MEMCHECK=1 turns it on)
  if [ "${STENCIL_CI_MEMCHECK:-0}" = 1 ]; then compute-sanitizer --tool memcheck ./build/bin/test_cuda "[pack]"; fi
  python -m pytest tests -q -m gpu
else
  echo "no CUDA device: GPU tier skipped"
fi
