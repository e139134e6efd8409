#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
run() { echo "## $*" >> $O; timeout 120 "$@" 2>/dev/null | grep device >> $O; }
for sz in "512 512 512" "1024 512 256" "1024 256 512" "512 1024 256" "512 256 1024" "2048 256 256" "256 1024 512"; do
  run $J $sz --exact-size -n 30 --tile-z 16 --device-time
done
run $J 1024 512 256 --exact-size -n 30 --tile-z 32 --device-time
run $J 1024 512 256 --exact-size -n 30 --tile-z 64 --device-time
run $J 1024 512 256 --exact-size -n 30 --tile-z 8 --device-time
STENCIL_JACOBI_DEBUG_SKIP=6 run $J 1024 512 256 --exact-size -n 30 --tile-z 16 --device-time
STENCIL_JACOBI_DEBUG_SKIP=6 run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time
run $J 1024 512 256 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
cat $O
