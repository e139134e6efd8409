#!/bin/bash
# Method sweep at 8 GPUs in one process (the counterpart of scripts/hal/run_weak_mp_method_*.sb): every legacy
# transport and the direct path, same problem.
set -euo pipefail
cd "$(dirname "$0")/../.."
export STENCIL_LOG_LEVEL=1
for m in "--direct --kernel" "--peer --kernel" "--staged" "--staged --peer --kernel --direct"; do
  ./build/bin/weak 512 512 512 30 $m
done
./build/bin/strong 1024 1024 1024 30
./build/bin/strong 1440 1452 700 30          # placement A/B of scripts/summit/run_1node_placement.sh
./build/bin/strong 1440 1452 700 30 --naive
