#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
N=${N:-4}
O=gpurun_out/exp_n$N.txt
: > $O
B=/root/repo/build/bin
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { echo "## $TAG $*" >> $O; ( cd /tmp && timeout 300 "$@" 2>/dev/null | grep -E "device|jacobi3d,|astaroth|verify|weak" ) >> $O; }
CUDA_VISIBLE_DEVICES=0 timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/pytest_gpu.log)" >> $O
timeout 600 $B/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n$N.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n$N.log)" >> $O
run $TR --master-port 29601 --no-python $B/jacobi3d 512 512 512 --verify 50 --tile-z 16 --delay-us 300
timeout 900 $TR --master-port 29621 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> $O
cat $O; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['verified'], d['exchange_us_per_iter'], d['astaroth']['ms_per_step_device'], d['e2e']['value'])"
