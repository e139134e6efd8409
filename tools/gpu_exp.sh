#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun). The kept measurement
# scripts are gpu_tests.sh (pytest -m gpu), gpu_multi.sh N (multi-GPU validation + benches), gpu_ncu.sh (ncu captures),
# gpu_perf.sh / gpu_apps.sh (app-level numbers).
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 500 ./build/bin/test_cuda "[jacobi]" 2>&1 | tail -1
timeout 300 python bench.py --steps 100 --warmup 10 2>/dev/null | tail -1 | cut -c1-200
