#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
PA=./build/prev_src/bld/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth|strong" >> $O; }
for g in "[astaroth]" "[jacobi]"; do
  timeout 900 ./build/bin/test_cuda "$g" > "gpurun_out/test_$g.log" 2>&1; echo "test $g rc=$? $(tail -1 "gpurun_out/test_$g.log")" >> $O
done
C="--x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20"
for rep in 1 2; do
TAG=prev run $PA $C
TAG=new run $A $C
TAG=new-nopush STENCIL_ASTAROTH_NO_FUSED_YZ=1 run $A $C
done
TAG=new run $A --x 323 --y 323 --z 162 --fields 8 --double --star19 --iters 20
TAG=new-tz54 run $A --x 323 --y 323 --z 162 --fields 8 --double --star19 --iters 20
run $A --x 100 --y 70 --z 90 --fields 3 --double --cross --iters 6 --check
run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time
run $J 645 645 323 --exact-size -n 30 --tile-z 16 --device-time
run $J 645 645 323 --exact-size -n 30 --tile-z 17 --device-time
run ./build/bin/strong 512 512 512 10 --compute --xplus 2 --xminus 1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -40; tail -4 gpurun_out/pytest_gpu.log
