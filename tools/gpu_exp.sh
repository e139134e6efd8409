#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth" >> $O; }
for g in "[astaroth]" "[jacobi]"; do
  timeout 900 ./build/bin/test_cuda "$g" > "gpurun_out/test_$g.log" 2>&1; echo "test $g rc=$? $(tail -1 "gpurun_out/test_$g.log")" >> $O
done
for cfg in "512 512 512" "645 645 323" "813 407 407" "1024 512 256"; do
  run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
done
C="--fields 8 --double --star19 --iters 20"
for tz in 32 52 64; do
  TAG=tz$tz run $A --x 256 --y 256 --z 256 $C --tile-z $tz
done
for tz in 32 54 41; do
  TAG=tz$tz run $A --x 323 --y 323 --z 162 $C --tile-z $tz
done
for tz in 32 51 41; do
  TAG=tz$tz run $A --x 406 --y 203 --z 203 $C --tile-z $tz
done
for tz in 32 64 43; do
  TAG=tz$tz run $A --x 512 --y 256 --z 128 $C --tile-z $tz
done
run $A --x 100 --y 70 --z 90 --fields 3 --double --cross --iters 6 --check
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -40
