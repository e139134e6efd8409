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
run() { echo "## $TAG $*" >> $O; ( cd /tmp && timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth|verify" ) >> $O; }
CUDA_VISIBLE_DEVICES=0 timeout 600 $B/test_cuda "[astaroth]" > gpurun_out/test_astaroth.log 2>&1; echo "test_astaroth rc=$? $(tail -1 gpurun_out/test_astaroth.log)" >> $O
timeout 600 $B/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n2.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n2.log)" >> $O
run $TR --master-port 29603 --no-python $B/astaroth-sim --x 96 --y 80 --z 72 --fields 3 --double --cross --iters 9 --check
TAG=n2 run $TR --master-port 29641 --no-python $B/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20 --weak
TAG=n2-1p run $B/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20 --weak
TAG=onegpu CUDA_VISIBLE_DEVICES=0 run $B/astaroth-sim --x 323 --y 323 --z 162 --fields 8 --double --star19 --iters 20
TAG=onegpu CUDA_VISIBLE_DEVICES=0 run $B/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
timeout 900 $TR --master-port 29621 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> $O
cat $O; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['verified'], d['exchange_us_per_iter'], d['astaroth'])"
