#!/bin/bash
cd /root/repo
export STENCIL_LOG_LEVEL=1
for sz in "645 645 323" "813 407 407" "1024 512 256"; do
A="./build/bin/jacobi3d $sz -n 40 --tile-z 16 --exact-size"
echo "== $sz wrap by load"; timeout 120 $A 2>/dev/null | tail -1
echo "== $sz x push"; STENCIL_JACOBI_NO_WRAP_LOAD=1 timeout 120 $A 2>/dev/null | tail -1
done
