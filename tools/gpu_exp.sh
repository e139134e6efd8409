#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 500 ./build/bin/test_cuda "[jacobi]" 2>&1 | grep "FAIL\|failed in\|test cases\|error" | head -8
timeout 300 python bench.py --steps 200 --warmup 10 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
