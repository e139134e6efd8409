#!/bin/bash
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
A="./build/bin/jacobi3d 512 512 512 -n 6 --tile-z 16"
$A > gpurun_out/ncu_plain_fused.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:jacobi7 -s 3 -c 1 -f -o gpurun_out/prof_fused $A > gpurun_out/ncu_fused.log 2>&1
echo "fused ncu rc=$?"
