#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
A=./build/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth" >> $O; }
timeout 900 ./build/bin/test_cuda "[astaroth]" > "gpurun_out/test_astaroth.log" 2>&1; echo "test rc=$? $(tail -1 gpurun_out/test_astaroth.log)" >> $O
C="--fields 8 --double --star19 --iters 20"
for tz in 32 43 52 64; do
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
cat $O
