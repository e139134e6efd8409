#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
timeout 900 ./build/bin/test_cuda "[async]" | tail -4
timeout 900 ./build/bin/test_cuda "[derivative]" | tail -2
