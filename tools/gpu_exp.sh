#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
run() { echo "## $PADS $SKEW $*" >> $O; STENCIL_ROW_PAD_UNITS=$PADS STENCIL_BUF_SKEW_BYTES=$SKEW timeout 300 "$@" 2>/dev/null | grep -E "device" >> $O; }
for cfg in "2048 256 256" "1024 512 256" "512 512 512"; do
  for PADS in 0 1 2 3 4 6 8 16; do
    SKEW=0 run $J $cfg --exact-size -n 30 --tile-z 16 --device-time --no-overlap
  done
  PADS=0
  for SKEW in 4096 65536 1048576 1310720; do
    run $J $cfg --exact-size -n 30 --tile-z 16 --device-time --no-overlap
  done
done
cat $O
