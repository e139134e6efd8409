#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth|verify|weak|^\[" >> $O; }
for g in "[jacobi]" "[astaroth]"; do
  timeout 900 ./build/bin/test_cuda "$g" > "gpurun_out/test_$g.log" 2>&1; echo "test $g rc=$? $(tail -1 "gpurun_out/test_$g.log")" >> $O
done
for cfg in "512 512 512" "1024 512 256" "645 645 323"; do
  TAG=zigzag run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
  TAG=nozz STENCIL_JACOBI_NO_ZIGZAG=1 run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
done
TAG=zigzag run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
TAG=nozz STENCIL_JACOBI_NO_ZIGZAG=1 run $J 512 512 512 --exact-size -n 30 --tile-z 16 --device-time --no-overlap
TAG=zz32 run $J 512 512 512 --exact-size -n 30 --tile-z 32 --device-time
TAG=zz8 run $J 512 512 512 --exact-size -n 30 --tile-z 8 --device-time
C="--x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20"
TAG=zigzag run $A $C
TAG=nozz STENCIL_ASTAROTH_NO_ZIGZAG=1 run $A $C
TAG=order1 STENCIL_ASTAROTH_TILE_ORDER=1 run $A $C
TAG=order2 STENCIL_ASTAROTH_TILE_ORDER=2 run $A $C
TAG=order2nozz STENCIL_ASTAROTH_TILE_ORDER=2 STENCIL_ASTAROTH_NO_ZIGZAG=1 run $A $C
TAG=cross run $A --x 256 --y 256 --z 256 --fields 8 --double --cross --iters 10
TAG=check run $A --x 100 --y 70 --z 90 --fields 3 --double --cross --iters 6 --check
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
ncu --metrics $M --clock-control none -k regex:jacobi7 -s 4 -c 2 --csv --log-file gpurun_out/ncu_jz.csv $J 512 512 512 --exact-size -n 8 --tile-z 16 > /dev/null 2>&1
echo "## ncu jacobi zigzag" >> $O; grep jacobi7 gpurun_out/ncu_jz.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 2 --csv --log-file gpurun_out/ncu_az.csv $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
echo "## ncu astaroth zigzag" >> $O; grep astaroth_tma gpurun_out/ncu_az.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -60
