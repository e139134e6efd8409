#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /root/repo
for mb in 2 3; do for tz in 16 32; do echo "minb=$mb tz=$tz"; STENCIL_JACOBI_MINB=$mb timeout 120 ./build/bin/jacobi3d 512 512 512 -n 50 --tile-z $tz 2>/dev/null; STENCIL_JACOBI_MINB=$mb timeout 120 ./build/bin/jacobi3d 512 512 512 -n 50 --tile-z $tz --no-overlap 2>/dev/null;  done; done
