#!/bin/bash
# scratch: new layer plan on N GPUs: protocol check with a delayed GPU, bench line
N=${1:-2}
mkdir -p gpurun_out
{
  echo "## verify with a delayed GPU, $N processes"
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 --no-python ./build/bin/jacobi3d 512 512 512 --verify 50 --delay-us 300 2>&1 | grep "jacobi3d-verify"
  echo "## bench"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --skip-exchange --skip-astaroth 2>gpurun_out/exp_bench.err | tee gpurun_out/exp_bench_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['ms_per_step_repeats'], d['value'], d['verified'], d['e2e']['value'])"
} > gpurun_out/exp.txt 2>&1
cat gpurun_out/exp.txt
