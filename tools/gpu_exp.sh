#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
N=${1:-2}
cd /tmp
for g in "[jacobi]"; do timeout 900 /root/repo/build/bin/test_cuda "$g" > /root/repo/gpurun_out/t.log 2>&1; echo "test $g rc=$? $(tail -1 /root/repo/gpurun_out/t.log)"; grep FAIL /root/repo/gpurun_out/t.log | head -5; done
echo "== 1 process x $N GPUs"; timeout 300 /root/repo/build/bin/jacobi3d 512 512 512 -n 50 --tile-z 16 2>/dev/null
echo "== 1 process x $N GPUs, exact even size"; timeout 300 /root/repo/build/bin/jacobi3d $((512*N)) 512 512 -n 50 --tile-z 16 --exact-size 2>/dev/null
echo "== $N processes"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 --no-python /root/repo/build/bin/jacobi3d 512 512 512 -n 50 --tile-z 16 2>/dev/null | grep jacobi3d
cd /root/repo
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 50 --warmup 5 --tile-z 16 2>/dev/null | cut -c1-200
