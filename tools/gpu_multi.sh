#!/bin/bash
# Multi-GPU validation: native tests on all visible GPUs (1 process), fused jacobi app, torchrun bench (1 process/GPU).
mkdir -p gpurun_out
N=${1:-2}
O=gpurun_out/multi.txt
: > $O
export STENCIL_LOG_LEVEL=2
(nvidia-smi topo -m || true) > gpurun_out/topo.txt 2>&1
timeout 900 ./build/bin/test_cuda "[multigpu]" > gpurun_out/test_multigpu.log 2>&1; echo "test_multigpu rc=$?" >> $O
timeout 900 ./build/bin/test_cuda "[exchange]" > gpurun_out/test_exchange.log 2>&1; echo "test_exchange rc=$?" >> $O
echo "1 process x $N GPUs, fused" >> $O
timeout 300 ./build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16 2>gpurun_out/j_1p.err >> $O
echo "1 process x $N GPUs, no-fused (exchange() schedule)" >> $O
timeout 300 ./build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16 --no-fused 2>>gpurun_out/j_1p.err >> $O
echo "$N processes x 1 GPU (torchrun --no-python jacobi3d), fused" >> $O
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 --no-python ./build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16 2>gpurun_out/j_np.err >> $O
echo "rc=$?" >> $O
echo "bench.py under torchrun, $N ranks" >> $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 30 --warmup 5 --tile-z 16 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> $O
echo "reference, $N GPUs" >> $O
timeout 900 python bench.py --impl reference --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?" >> $O
cat $O; tail -5 gpurun_out/test_multigpu.log; tail -3 gpurun_out/test_exchange.log; cat gpurun_out/bench_n$N.json gpurun_out/bench_ref_n$N.json; tail -5 gpurun_out/j_np.err gpurun_out/bench_n$N.err
