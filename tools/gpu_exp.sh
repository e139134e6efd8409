#!/bin/bash
mkdir -p gpurun_out
export STENCIL_LOG_LEVEL=1
cd /root/repo
timeout 900 ./build/bin/test_cuda > gpurun_out/t.log 2>&1; echo "test_cuda rc=$? $(tail -1 gpurun_out/t.log)"; grep FAIL gpurun_out/t.log | head
echo "weak 512 r3 x4"; timeout 120 ./build/bin/weak 512 512 512 30 2>/dev/null | tail -2
echo "bench-pack"; timeout 120 ./build/bin/bench-pack 2>/dev/null | tail -8
echo "bench-exchange"; timeout 120 ./build/bin/bench-exchange 2>/dev/null | tail -4
echo "astaroth"; timeout 120 ./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star --iters 10 2>/dev/null
echo "jacobi"; timeout 120 ./build/bin/jacobi3d 512 512 512 -n 30 --tile-z 16 2>/dev/null | tail -1
