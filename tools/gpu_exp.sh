#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /root/repo
timeout 900 ./build/bin/test_cuda "[jacobi]" > gpurun_out/t.log 2>&1; echo "test jacobi rc=$? $(tail -1 gpurun_out/t.log)"
for mb in 2 3; do for tz in 16 32 64; do echo "minb=$mb tz=$tz"; STENCIL_JACOBI_MINB=$mb timeout 120 ./build/bin/jacobi3d 512 512 512 -n 50 --tile-z $tz 2>/dev/null; STENCIL_JACOBI_MINB=$mb timeout 120 ./build/bin/jacobi3d 512 512 512 -n 50 --tile-z $tz --no-overlap 2>/dev/null;  done; done
timeout 600 python bench.py --steps 50 --warmup 5 --tile-z 32 2>/dev/null | cut -c1-330
