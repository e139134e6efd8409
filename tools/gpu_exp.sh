#!/bin/bash
mkdir -p gpurun_out
cd /root/repo
timeout 900 python -m pytest tests/test_python_gpu.py tests/test_native_gpu.py -m gpu -x -q 2>&1 | tail -5
for tz in 16 32; do echo "tile-z $tz"; timeout 300 python bench.py --steps 200 --warmup 10 --tile-z $tz 2>/dev/null | tail -1 | cut -c1-200; done
timeout 300 python bench.py --steps 200 --warmup 10 --tile-z 16 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['clocks'], d['e2e'])"
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['clocks'])"
