#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /root/repo
timeout 900 ./build/bin/test_cuda "[astaroth]" > gpurun_out/t.log 2>&1; echo "test astaroth rc=$? $(tail -1 gpurun_out/t.log)"; grep FAIL gpurun_out/t.log | head
A="./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 10"
echo "tma overlap"; timeout 120 $A 2>/dev/null
echo "tma no-overlap"; timeout 120 $A --no-overlap 2>/dev/null
echo "ldg overlap"; STENCIL_ASTAROTH_NO_TMA=1 timeout 120 $A 2>/dev/null
echo "fp32 x4 512^3 tma no-overlap"; timeout 120 ./build/bin/astaroth-sim --x 512 --y 512 --z 512 --fields 4 --star --iters 10 --no-overlap 2>/dev/null
