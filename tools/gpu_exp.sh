#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
OLD=./build/old_src/bld/bin/astaroth-sim
run() { echo "## $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth|verify" >> $O; }
timeout 600 ./build/bin/test_cuda "[jacobi]" > gpurun_out/test_jacobi.log 2>&1; echo "test_jacobi rc=$? $(tail -1 gpurun_out/test_jacobi.log)" >> $O
timeout 600 ./build/bin/test_cuda "[tx]" > gpurun_out/test_tx.log 2>&1; echo "test_tx rc=$? $(tail -1 gpurun_out/test_tx.log)" >> $O
for cfg in "512 512 512" "1024 512 256" "2048 256 256" "645 645 323" "813 407 407"; do
  run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
done
STENCIL_JACOBI_DEBUG_SKIP=8 run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time
STENCIL_JACOBI_DEBUG_SKIP=8 run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
run $J 300 200 100 --exact-size --verify 20 --tile-z 16 --delay-us 100
echo "## xface microbench" >> $O
./build/xface 4 >> $O 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 24 -c 24 --csv --log-file gpurun_out/ncu_xface.csv ./build/xface 4 > /dev/null 2>&1
grep -E "k_scalar|k_vec|k_tma" gpurun_out/ncu_xface.csv | awk -F'","' '{print $5, $(NF-2), $(NF)}' | sort | uniq -c | sort -k2 >> $O
for i in 1 2; do
run $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
run $OLD --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
STENCIL_ASTAROTH_NO_FUSED_YZ=1 run $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
STENCIL_ASTAROTH_NO_FUSED_YZ=1 run $OLD --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
done
M=gpu__time_duration.sum
STENCIL_ASTAROTH_NO_FUSED_YZ=1 ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 2 --csv --log-file gpurun_out/ncu_ast_new.csv $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
STENCIL_ASTAROTH_NO_FUSED_YZ=1 ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 2 --csv --log-file gpurun_out/ncu_ast_old.csv $OLD --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
echo "## ncu astaroth new no-fused-yz" >> $O; grep astaroth_tma gpurun_out/ncu_ast_new.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
echo "## ncu astaroth old no-fused-yz" >> $O; grep astaroth_tma gpurun_out/ncu_ast_old.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
cat $O
