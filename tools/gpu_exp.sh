#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun). The kept measurement
# scripts are gpu_tests.sh (pytest -m gpu), gpu_multi.sh N (multi-GPU validation + benches), gpu_ncu.sh (ncu captures),
# gpu_perf.sh / gpu_apps.sh (app-level numbers).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
timeout 600 ./build/bin/test_cuda "[jacobi]" > gpurun_out/test_jacobi.log 2>&1; echo "test_jacobi rc=$? $(tail -1 gpurun_out/test_jacobi.log)" >> $O
J=./build/bin/jacobi3d
for sz in "512 512 512" "645 645 323" "813 407 407" "1024 512 256"; do
  echo "## $sz narrow" >> $O; timeout 120 $J $sz --exact-size -n 30 --tile-z 16 --device-time 2>/dev/null >> $O
  echo "## $sz wide only" >> $O; STENCIL_JACOBI_NO_NARROW=1 timeout 120 $J $sz --exact-size -n 30 --tile-z 16 --device-time 2>/dev/null >> $O
done
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> $O
cat $O; grep -B2 -A8 FAIL gpurun_out/test_jacobi.log | head -40; cat gpurun_out/bench1.json
