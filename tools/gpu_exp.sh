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
run() { echo "## $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth|verify|weak|^\[" >> $O; }
for g in "[pack]" "[translate]" "[packer]" "[exchange]" "[jacobi]" "[tx]"; do
  timeout 900 ./build/bin/test_cuda "$g" > "gpurun_out/test_$g.log" 2>&1; echo "test $g rc=$? $(tail -1 "gpurun_out/test_$g.log")" >> $O
done
for cfg in "512 512 512" "1024 512 256" "2048 256 256"; do
  run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
done
run ./build/bin/bench-pack --iters 30
run ./baseline/_ref/bin/bench-pack --iters 30
run ./build/bin/weak 512 512 512 30
run ./build/bin/weak-exchange 512 512 512 30
STENCIL_EXCHANGE_TRACE=2 ./build/bin/weak 512 512 512 4 2>&1 | grep "exchange " | tail -2 >> $O
export STENCIL_ASTAROTH_NO_FUSED_YZ=1
C="--x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6"
ncu --set full --clock-control none --import-source on -k regex:astaroth_tma -s 3 -c 1 -f -o gpurun_out/prof_ast_new $A $C > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:astaroth_tma -s 3 -c 1 -f -o gpurun_out/prof_ast_old $OLD $C > /dev/null 2>&1
unset STENCIL_ASTAROTH_NO_FUSED_YZ
cat $O; grep -h -B2 -A8 FAIL gpurun_out/test_*.log | head -60
