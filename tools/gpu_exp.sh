#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /tmp
for g in "[exchange]" "[pack]" "[packer]" "[astaroth]" "[jacobi]"; do timeout 900 /root/repo/build/bin/test_cuda "$g" > /root/repo/gpurun_out/t.log 2>&1; echo "test $g rc=$? $(tail -1 /root/repo/gpurun_out/t.log)"; grep FAIL /root/repo/gpurun_out/t.log | head -5; done
echo "== 1 process x 2 GPUs, weak 512^3/GPU"; STENCIL_EXCHANGE_TRACE=2 timeout 300 /root/repo/build/bin/weak 512 512 512 30 2>&1 | grep -E "^exchange|^weak"
echo "== exact 1024x512x512 (even sizes)"; STENCIL_EXCHANGE_TRACE=2 timeout 300 /root/repo/build/bin/strong 1024 512 512 30 2>&1 | grep -E "^exchange|^weak"
echo "== 2 processes"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 --no-python /root/repo/build/bin/strong 1024 512 512 30 2>&1 | grep -E "^weak"
echo "== reference peer+kernel"; timeout 300 /root/repo/baseline/_ref/bin/strong 1024 512 512 30 --peer --kernel 2>/dev/null | grep -E "^weak"
timeout 300 /root/repo/build/bin/bench-pack
