#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
N=2
B=/root/repo/build/bin
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { echo "## $TAG $*" >> $O; ( cd /tmp && timeout 300 "$@" 2>/dev/null | grep -E "device|weak" ) >> $O; }
timeout 600 $B/test_cuda "[exchange]" > gpurun_out/test_exchange_n2.log 2>&1; echo "test_exchange rc=$? $(tail -1 gpurun_out/test_exchange_n2.log)" >> $O
timeout 600 $B/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n2.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n2.log)" >> $O
run $B/weak-exchange 512 512 512 30
run $TR --master-port 29631 --no-python $B/weak-exchange 512 512 512 30
CUDA_VISIBLE_DEVICES=0 run $B/weak-exchange 512 512 512 30
STENCIL_EXCHANGE_TRACE=3 $B/weak 512 512 512 4 2>&1 | grep "exchange " | tail -4 >> $O
cat $O
