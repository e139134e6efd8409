#!/bin/bash
# ncu --set full captures of the hot kernels on one GPU (each only after the same command has run cleanly without ncu);
# reports land in gpurun_out/, condensed afterwards with tools/ncu_summary.py into profiles/ncu_*.csv.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export STENCIL_LOG_LEVEL=1
cap() { # name, kernel regex, skip, command...   (ONLY="name name ..." restricts the captures)
  local name=$1 rx=$2 skip=$3; shift 3
  if [ -n "$ONLY" ] && ! echo " $ONLY " | grep -q " $name "; then return; fi
  "$@" > gpurun_out/ncu_plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o gpurun_out/prof_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name ncu rc=$?"
}
J="./build/bin/jacobi3d 512 512 512 --exact-size -n 6 ${TZ:+--tile-z $TZ}"
cap jacobi_fused jacobi7 3 $J
cap jacobi_plain jacobi7 3 $J --no-overlap
cap jacobi_narrow jacobi7 3 ./build/bin/jacobi3d 645 645 323 --exact-size -n 6 ${TZ:+--tile-z $TZ}
cap astaroth astaroth_tma 2 ./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 4
cap astaroth_cross astaroth_tma 2 ./build/bin/astaroth-sim --x 256 --y 256 --z 256 --fields 8 --double --cross --iters 4
cap boxcopy box_copy 2 ./build/bin/weak 512 512 512 4
