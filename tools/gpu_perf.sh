#!/bin/bash
# Jacobi kernel performance sweep on one GPU: kernel variant x tile depth, through the jacobi3d app and bench.py.
mkdir -p gpurun_out
O=gpurun_out/perf.txt
: > $O
export STENCIL_LOG_LEVEL=1
timeout 600 ./build/bin/test_cuda "[jacobi]" > gpurun_out/test_jacobi.log 2>&1; echo "test_jacobi rc=$?" >> $O
for tz in 8 16 32 64; do
  echo "tma fused tileZ=$tz" >> $O; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 30 --tile-z $tz 2>/dev/null >> $O
done
for tz in 16 32 64; do
echo "tma no-overlap tileZ=$tz" >> $O; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 30 --no-overlap --tile-z $tz 2>/dev/null >> $O
done
echo "tma no-fused" >> $O; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 30 --no-fused 2>/dev/null >> $O
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench rc=$?" >> $O
cat $O; tail -3 gpurun_out/test_jacobi.log; cat gpurun_out/bench2.json
