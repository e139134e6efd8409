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
STENCIL_JACOBI_PDL=1 timeout 900 ./build/bin/test_cuda "[jacobi]" > "gpurun_out/test_jacobi_pdl.log" 2>&1; echo "test pdl rc=$? $(tail -1 gpurun_out/test_jacobi_pdl.log)" >> $O
for rep in 1 2; do
for cfg in "512 512 512" "645 645 323"; do
  TAG=nopdl run $J $cfg --exact-size -n 40 --tile-z 16 --device-time
  TAG=pdl STENCIL_JACOBI_PDL=1 run $J $cfg --exact-size -n 40 --tile-z 16 --device-time
done
done
TAG=pdl STENCIL_JACOBI_PDL=1 run $J 300 200 100 --exact-size --verify 30 --tile-z 16 --delay-us 50
TAG=pdl-nooverlap STENCIL_JACOBI_PDL=1 run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
TAG=nopdl-nooverlap run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
timeout 300 python bench.py --steps 30 --warmup 5 --random-init --skip-exchange --skip-astaroth > gpurun_out/bench1_random.json 2> gpurun_out/bench1_random.err; echo "bench random rc=$?" >> $O
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> $O 2>&1
cat $O; cut -c1-400 gpurun_out/bench1_random.json; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -30
