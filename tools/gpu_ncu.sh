#!/bin/bash
# ncu capture of the Jacobi kernels (fused and plain) on one GPU; reports land in gpurun_out/.
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
A="./build/bin/jacobi3d 512 512 512 -n 6 --tile-z ${TZ:-16}"
$A > gpurun_out/ncu_plain_fused.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:jacobi7 -s 3 -c 1 -f -o gpurun_out/prof_fused $A > gpurun_out/ncu_fused.log 2>&1
echo "fused ncu rc=$?"
$A --no-overlap > gpurun_out/ncu_plain_noov.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:jacobi7 -s 3 -c 1 -f -o gpurun_out/prof_plain $A --no-overlap > gpurun_out/ncu_noov.log 2>&1
echo "plain ncu rc=$?"
B="./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 4"
$B > gpurun_out/ncu_plain_astaroth.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:astaroth_tma -s 2 -c 1 -f -o gpurun_out/prof_astaroth $B > gpurun_out/ncu_astaroth.log 2>&1
echo "astaroth ncu rc=$?"
C="./build/bin/weak 512 512 512 4"
$C > gpurun_out/ncu_plain_weak.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:box_copy -s 2 -c 1 -f -o gpurun_out/prof_boxcopy $C > gpurun_out/ncu_weak.log 2>&1
echo "box_copy ncu rc=$?"
