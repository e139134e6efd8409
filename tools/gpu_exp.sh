#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 900 ./build/bin/test_cuda "[exchange]" | tail -1
timeout 900 ./build/bin/test_cuda "[astaroth]" | tail -1
timeout 900 ./build/bin/test_cuda "[jacobi]" | tail -1
timeout 900 ./build/bin/test_cuda "[pack]" | tail -1
STENCIL_EXCHANGE_TRACE=5 timeout 120 ./build/bin/weak 512 512 512 6 2>&1 | grep "^exchange" | tail -1
timeout 120 ./build/bin/weak 512 512 512 30 2>/dev/null | tail -1 | cut -d, -f 20-
timeout 120 ./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 20 2>/dev/null | tail -1
