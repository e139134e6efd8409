#!/bin/bash
# One GPU-box pass: native + python GPU tests, headline bench, the jacobi3d app. Results land in gpurun_out/.
mkdir -p gpurun_out
S=gpurun_out/summary.txt
: > $S
(nvidia-smi -L || true) > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> $S
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> $S
for tz in 8 16 32 64 128; do
  timeout 120 ./build/bin/jacobi3d 512 512 512 -n 20 --tile-z $tz 2>> gpurun_out/jacobi3d.err >> gpurun_out/jacobi3d.csv
done
timeout 120 ./build/bin/jacobi3d 512 512 512 -n 20 --no-fused 2>> gpurun_out/jacobi3d.err >> gpurun_out/jacobi3d.csv
timeout 120 ./build/bin/jacobi3d 512 512 512 -n 20 --no-overlap 2>> gpurun_out/jacobi3d.err >> gpurun_out/jacobi3d.csv
echo "apps done" >> $S
cat $S
tail -n 30 gpurun_out/pytest_gpu.log
cat gpurun_out/bench1.json
cat gpurun_out/jacobi3d.csv
