#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 900 ./build/bin/test_cuda "[jacobi]" | tail -1
A="./build/bin/jacobi3d 512 512 512 -n 40 --tile-z 16"
echo "epilogue x+yz"; timeout 120 $A 2>/dev/null | tail -1
echo "skip x"; STENCIL_JACOBI_DEBUG_SKIP=2 timeout 120 $A 2>/dev/null | tail -1
echo "skip all"; STENCIL_JACOBI_DEBUG_SKIP=7 timeout 120 $A 2>/dev/null | tail -1
timeout 300 python bench.py --steps 200 --warmup 10 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
