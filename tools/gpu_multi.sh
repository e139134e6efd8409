#!/bin/bash
# Multi-GPU validation: native tests on all visible GPUs (1 process), fused jacobi app, torchrun bench (1 process/GPU).
mkdir -p gpurun_out
N=${1:-2}
O=gpurun_out/multi_n$N.txt
: > $O
export STENCIL_LOG_LEVEL=1
R=baseline/_ref/bin
ROOT=$(pwd)
(nvidia-smi topo -m || true) > gpurun_out/topo_n$N.txt 2>&1
timeout 900 ./build/bin/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n$N.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n$N.log)" >> $O
run() { echo "## $*" >> $O; ( cd /tmp && timeout 600 "$@" 2>/dev/null | tail -6 ) >> $O; }
run $ROOT/build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16
run python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 --no-python $ROOT/build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16
run $ROOT/build/bin/weak 512 512 512 30
[ -n "$SKIP_REF" ] || run $ROOT/$R/weak 512 512 512 30 --peer --kernel
run python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 --no-python $ROOT/build/bin/weak 512 512 512 30
run $ROOT/build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 10 --weak
echo "## bench.py under torchrun, $N ranks" >> $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 30 --warmup 5 --tile-z 16 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> $O
echo "## reference, $N GPUs" >> $O
[ -n "$SKIP_REF" ] || timeout 1200 python bench.py --impl reference --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?" >> $O
cat $O; cat gpurun_out/bench_n$N.json gpurun_out/bench_ref_n$N.json | cut -c1-420; tail -n 3 gpurun_out/bench_n$N.err
