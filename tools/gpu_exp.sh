#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
for g in "[jacobi]" "[astaroth]"; do
  timeout 900 ./build/bin/test_cuda "$g" > "gpurun_out/test_$g.log" 2>&1; echo "test $g rc=$? $(tail -1 "gpurun_out/test_$g.log")" >> $O
done
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
run() { echo "## $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth" >> $O; }
run $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
run $A --x 323 --y 323 --z 162 --fields 8 --double --star19 --iters 20
STENCIL_ASTAROTH_NO_NARROW=1 run $A --x 323 --y 323 --z 162 --fields 8 --double --star19 --iters 20
run $A --x 406 --y 203 --z 203 --fields 8 --double --star19 --iters 20
run $A --x 512 --y 256 --z 128 --fields 8 --double --star19 --iters 20
run $A --x 256 --y 256 --z 256 --fields 8 --double --cross --iters 10
run $A --x 100 --y 70 --z 50 --fields 3 --double --cross --iters 6 --check
for sz in "512 512 512" "1024 512 256" "1024 256 512" "512 1024 256" "2048 256 256" "256 1024 512"; do
  run $J $sz --exact-size -n 30 --tile-z 16 --device-time
done
run $J 1024 512 256 --exact-size -n 30 --tile-z 32 --device-time
run $J 1024 512 256 --exact-size -n 30 --tile-z 8 --device-time
STENCIL_JACOBI_DEBUG_SKIP=6 run $J 1024 512 256 --exact-size -n 30 --tile-z 16 --device-time
run $J 1024 512 256 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> $O
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?" >> $O
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -60; cat gpurun_out/bench1.json gpurun_out/bench_ref.json; tail -5 gpurun_out/bench1.err; tail -8 gpurun_out/pytest_gpu.log
