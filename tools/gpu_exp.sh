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
run() { echo "## $TAG $*" >> $O; ( cd /tmp && timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth-sim|verify" ) >> $O; }
timeout 600 $B/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n2.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n2.log)" >> $O
TAG=default run $TR --master-port 29611 --no-python $B/jacobi3d 512 512 512 -n 40 --tile-z 16 --device-time
TAG=natural STENCIL_JACOBI_NATURAL_ORDER=1 run $TR --master-port 29612 --no-python $B/jacobi3d 512 512 512 -n 40 --tile-z 16 --device-time


TAG=onegpu CUDA_VISIBLE_DEVICES=0 run $B/jacobi3d 645 645 323 --exact-size -n 40 --tile-z 16 --device-time
TAG=default run $TR --master-port 29614 --no-python $B/jacobi3d 512 512 512 -n 40 --tile-z 16 --device-time

cat $O
