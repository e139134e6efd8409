#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun). The kept measurement
# scripts are gpu_tests.sh (pytest -m gpu), gpu_multi.sh N (multi-GPU validation + benches), gpu_ncu.sh (ncu captures),
# gpu_perf.sh / gpu_apps.sh (app-level numbers).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
N=${N:-4}
timeout 120 ./build/bin/test_cuda "[multigpu]" 2>&1 | tail -1
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 100 --warmup 5 2>/dev/null | tail -1 > gpurun_out/bench_n$N.json
python -c "import json; d=json.loads(open('gpurun_out/bench_n$N.json').read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
