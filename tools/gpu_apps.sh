#!/bin/bash
# Apps pass on N GPUs (default: all visible): astaroth tests, exchange benchmarks for ours and the reference.
mkdir -p gpurun_out
O=gpurun_out/apps.txt
: > $O
export STENCIL_LOG_LEVEL=1
R=baseline/_ref/bin
cd $(dirname $0)/.. ; ROOT=$(pwd)
run() { echo "## $*" >> $O; ( cd /tmp && timeout 300 "$@" 2>/dev/null | tail -12 ) >> $O; }
for g in "[astaroth]" "[pack]" "[translate]" "[packer]" "[exchange]" "[jacobi]"; do
timeout 900 ./build/bin/test_cuda "$g" > gpurun_out/test_$g.log 2>&1; echo "test $g rc=$? $(tail -1 gpurun_out/test_$g.log)" >> $O
done
run $ROOT/build/bin/bench-pack --iters 30
run $ROOT/$R/bench-pack --iters 30
run $ROOT/build/bin/weak 512 512 512 30
run $ROOT/$R/weak 512 512 512 30 --peer --kernel
run $ROOT/build/bin/weak-exchange 512 512 512 30
run $ROOT/$R/weak-exchange 512 512 512 30 --peer --kernel
run $ROOT/build/bin/bench-exchange --x 256 --y 256 --z 256 --q 4 --fr 3 --er 3
cat $O; grep -h FAIL gpurun_out/test_*.log | head
