#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|verify" >> $O; }
timeout 900 ./build/bin/test_cuda "[jacobi]" > "gpurun_out/test_jacobi.log" 2>&1; echo "test rc=$? $(tail -1 gpurun_out/test_jacobi.log)" >> $O
for rep in 1 2; do
for cfg in "512 512 512" "1024 512 256"; do
  TAG=lean run $J $cfg --exact-size -n 40 --tile-z 16 --device-time
  TAG=fat STENCIL_JACOBI_FAT_KERNEL=1 run $J $cfg --exact-size -n 40 --tile-z 16 --device-time
done
done
TAG=lean run $J 512 512 512 --exact-size -n 40 --tile-z 16 --device-time --no-overlap
TAG=fat STENCIL_JACOBI_FAT_KERNEL=1 run $J 512 512 512 --exact-size -n 40 --tile-z 16 --device-time --no-overlap
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -30
