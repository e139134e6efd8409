#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 500 ./build/bin/test_cuda "[multigpu]" 2>&1 | tail -1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 2 --steps 100 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
STENCIL_PARTITION_TIES=x timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 2 --steps 100 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('x-cut', d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 200 python -m pytest tests/test_multiprocess_gpu.py -m gpu -x -q 2>&1 | tail -1
