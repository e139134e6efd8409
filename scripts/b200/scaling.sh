#!/bin/bash
# Weak-scaling campaign on one 8 x B200 box (the counterpart of the reference's scripts/summit/*_weak_*.sh and
# scripts/hal/run_weak_*.sb): jacobi3d, weak exchange and astaroth-sim at 1, 2, 4, 8 GPUs, one process per GPU.
# Usage: scripts/b200/scaling.sh [out_dir]
set -euo pipefail
cd "$(dirname "$0")/../.."
OUT=${1:-scaling_out}
mkdir -p "$OUT"
export STENCIL_LOG_LEVEL=1
for n in 1 2 4 8; do
  launch="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) --no-python"
  $launch ./build/bin/jacobi3d 512 512 512 -n 100 | tee -a "$OUT/jacobi3d.csv"
  $launch ./build/bin/weak 512 512 512 30 | tee -a "$OUT/weak.csv"
  # through a wrapper: torchrun's own parser takes "--star" for an abbreviation of --start-method
  EXTRA=--weak ITERS=20 $launch tools/astaroth_target.sh | tee -a "$OUT/astaroth.csv"
done
