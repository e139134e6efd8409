#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 900 ./build/bin/test_cuda "[astaroth]" | tail -2
timeout 120 ./build/bin/astaroth-sim --x 323 --y 323 --z 323 --fields 8 --double --star --iters 10 2>/dev/null | tail -1
STENCIL_ASTAROTH_NO_FUSED_X=1 timeout 120 ./build/bin/astaroth-sim --x 323 --y 323 --z 323 --fields 8 --double --star --iters 10 2>/dev/null | tail -1
