#!/bin/bash
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
N=8
timeout 600 ./build/bin/test_cuda "[multigpu]" > gpurun_out/test_multigpu_n8b.log 2>&1; echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n8b.log)"
for t in z x; do
  echo "== ties $t"
  STENCIL_PARTITION_TIES=$t timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/bench_n8_ties_$t.json 2> gpurun_out/bench_n8_ties_$t.err; echo "rc=$?"; cut -c1-170 gpurun_out/bench_n8_ties_$t.json
done
echo "== weak ties z"; ( cd /tmp && timeout 300 /root/repo/build/bin/weak 512 512 512 30 2>/dev/null | tail -1 | cut -d, -f 2-4,18-22 )
echo "== astaroth ties z"; ( cd /tmp && timeout 300 /root/repo/build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 20 --weak 2>/dev/null | tail -1 )
