#!/bin/bash
# Multi-GPU validation on N GPUs of one box: native [multigpu] tests (1 process), large fused-vs-plain verification with
# one GPU held back (1 process and 1 process per GPU), apps in both deployments, bench.py under torchrun, reference arm.
#   tools/gpu_multi.sh N        (SKIP_REF=1 leaves the reference's runs out)
mkdir -p gpurun_out
N=${1:-2}
O=gpurun_out/multi_n$N.txt
: > $O
export STENCIL_LOG_LEVEL=1
R=baseline/_ref/bin
ROOT=$(pwd)
B=$ROOT/build/bin
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(nvidia-smi topo -m || true) > gpurun_out/topo_n$N.txt 2>&1
timeout 900 $B/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n$N.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n$N.log)" >> $O
run() { echo "## $*" >> $O; ( cd /tmp && timeout 600 "$@" 2>/dev/null | grep -v "^$" | tail -6 ) >> $O; }
# correctness at the bench size: 50 iterations, fused == plain bit for bit, rank 0's GPU held back 300 us per iteration
run $B/jacobi3d 512 512 512 --verify 50 --tile-z 16 --delay-us 300
run $TR --master-port 29601 --no-python $B/jacobi3d 512 512 512 --verify 50 --tile-z 16 --delay-us 300
[ -n "$QUICK" ] || run $TR --master-port 29602 --no-python $B/jacobi3d 512 512 512 --verify 50 --tile-z 16 --delay-us 300 --delay-rank 1
run $TR --master-port 29603 --no-python $B/astaroth-sim --x 96 --y 80 --z 72 --fields 3 --double --cross --iters 9 --check
run $B/astaroth-sim --x 96 --y 80 --z 72 --fields 3 --double --cross --iters 9 --check
# timing, both deployments
run $B/jacobi3d 512 512 512 -n 30 --tile-z 16 --device-time
run $TR --master-port 29611 --no-python $B/jacobi3d 512 512 512 -n 30 --tile-z 16 --device-time
run $B/weak 512 512 512 30
run $TR --master-port 29631 --no-python $B/weak 512 512 512 30
[ -n "$SKIP_REF" ] || run $ROOT/$R/weak 512 512 512 30 --peer --kernel
run $B/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20 --weak
run $TR --master-port 29641 --no-python $B/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20 --weak
# NVLink bytes of the fused kernel (1 process x N GPUs under ncu; kernels are serialised by the profiler, flags refer to the
# previous iteration, so nothing can wait for ever; bounded by timeout all the same)
NV=gpu__time_duration.sum,nvltx__bytes.sum,nvlrx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_user.sum,dram__bytes_read.sum,dram__bytes_write.sum
( cd /tmp && timeout 300 ncu --metrics $NV --clock-control none -k regex:jacobi7 -s $((4 * N)) -c $((2 * N)) --csv --log-file $ROOT/gpurun_out/ncu_nvlink_n$N.csv \
    $B/jacobi3d 512 512 512 -n 8 --tile-z 16 > /dev/null 2>&1 ); echo "ncu nvlink rc=$?" >> $O
grep jacobi7 gpurun_out/ncu_nvlink_n$N.csv 2>/dev/null | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
echo "## bench.py under torchrun, $N ranks" >> $O
timeout 900 $TR --master-port 29621 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?" >> $O
echo "## reference, $N GPUs" >> $O
[ -n "$SKIP_REF" ] || timeout 1200 python bench.py --impl reference --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?" >> $O
cat $O; cat gpurun_out/bench_n$N.json gpurun_out/bench_ref_n$N.json 2>/dev/null | cut -c1-1500; tail -n 3 gpurun_out/bench_n$N.err
