#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 900 ./build/bin/test_cuda "[astaroth]" | tail -6
timeout 900 ./build/bin/test_cuda "[exchange]" | tail -1
A="./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 20"
echo "fused x"; timeout 120 $A 2>/dev/null | tail -1
echo "no fused x"; STENCIL_ASTAROTH_NO_FUSED_X=1 timeout 120 $A 2>/dev/null | tail -1
echo "fp32 x4 512^3"; timeout 120 ./build/bin/astaroth-sim --x 512 --y 512 --z 512 --fields 4 --star --iters 10 2>/dev/null | tail -1
echo "fp32 x4 512^3 no fused x"; STENCIL_ASTAROTH_NO_FUSED_X=1 timeout 120 ./build/bin/astaroth-sim --x 512 --y 512 --z 512 --fields 4 --star --iters 10 2>/dev/null | tail -1
