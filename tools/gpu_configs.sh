#!/bin/bash
# BASELINE configs 4 and 5 and the box's link matrix on N GPUs of one box (both arms where the reference can run it):
#   strong scaling 1024^3 (exchange-only: both arms; with uneven x radius + overlapped compute: own arm),
#   bench-exchange size sweep (both arms), measure-buf-exchange and bench-alltoallv (own: memcpy arm vs direct-store arm).
#   tools/gpu_configs.sh N [quick]
mkdir -p gpurun_out
N=${1:-8}
O=gpurun_out/configs_n$N.txt
: > $O
export STENCIL_LOG_LEVEL=1
ROOT=$(pwd)
B=$ROOT/build/bin
R=$ROOT/baseline/_ref/bin
run() { echo "## $*" >> $O; ( cd /tmp && timeout 900 "$@" 2>/dev/null | grep -v "^$" | tail -${TAIL:-4} ) >> $O; }
# config 4: strong scaling, whole domain 1024^3 (and an uneven one), 4 x fp32, radius 3
run $B/strong 1024 1024 1024 20
run $R/strong 1024 1024 1024 20 --peer --kernel
run $B/strong 1000 1023 1024 20
run $B/strong 1024 1024 1024 20 --xplus 2 --xminus 1 --compute
run $B/strong 1000 1023 1024 20 --xplus 2 --xminus 1 --compute
# config 5: bench-exchange message-size sweep, 4 quantities, radius 3 everywhere
for s in 64 128 256 512 1024; do
  TAIL=7 run $B/bench-exchange --x $s --y $s --z $s --q 4 --fr 3 --er 3 --cr 3 --iters 20
  [ -n "$SKIP_REF" ] || TAIL=7 run $R/bench-exchange --x $s --y $s --z $s --q 4 --fr 3 --er 3 --cr 3 --iters 20
done
# the box's link matrix under full contention, and the stencil-like all-to-all with copies vs direct stores
TAIL=40 run $B/measure-buf-exchange --gpus $N --rounds 12 --measures 5
TAIL=40 run $B/bench-alltoallv --gpus $N --iters 10
cat $O
