#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
bash tools/gpu_ncu.sh > gpurun_out/ncu_all.log 2>&1
SAN_TIMEOUT=150 SAN_GROUPS="[pack] [exchange] [jacobi]" bash tools/gpu_sanitize.sh > gpurun_out/sanitize.log 2>&1
cat gpurun_out/sanitizer/summary.txt gpurun_out/ncu_all.log
