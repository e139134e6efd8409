#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /root/repo
A="./build/bin/jacobi3d 512 512 512 -n 40 --tile-z 16"
for s in 0 2 4 6 7; do echo "skip=$s"; STENCIL_JACOBI_DEBUG_SKIP=$s timeout 120 $A 2>/dev/null | tail -1; done
echo "no-fused (overlap)"; timeout 120 $A --no-fused 2>/dev/null | tail -1
echo "no-overlap"; timeout 120 $A --no-fused --no-overlap 2>/dev/null | tail -1
echo "tile-z 32"; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 40 --tile-z 32 2>/dev/null | tail -1
echo "tile-z 8"; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 40 --tile-z 8 2>/dev/null | tail -1
