#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/exp.txt
: > $O
export STENCIL_LOG_LEVEL=1
A="./build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16"
for sk in 0 1 2 4 6 7; do
  echo "skip=$sk fused" >> $O; STENCIL_JACOBI_DEBUG_SKIP=$sk timeout 120 $A 2>/dev/null >> $O
done
echo "no-overlap" >> $O;  timeout 120 $A --no-overlap 2>/dev/null >> $O
cat $O
