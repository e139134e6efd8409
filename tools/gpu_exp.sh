#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
J=./build/bin/jacobi3d
A=./build/bin/astaroth-sim
PJ=./build/prev_src/bld/bin/jacobi3d
PA=./build/prev_src/bld/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth-sim" >> $O; }
for rep in 1 2; do
for cfg in "512 512 512" "1024 512 256"; do
  TAG=prev run $PJ $cfg --exact-size -n 30 --tile-z 16 --device-time
  TAG=new run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
  TAG=new-nozz STENCIL_JACOBI_NO_ZIGZAG=1 run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
  TAG=new-nokeep STENCIL_NO_L2_KEEP=1 run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
  TAG=new-nozz-nokeep STENCIL_NO_L2_KEEP=1 STENCIL_JACOBI_NO_ZIGZAG=1 run $J $cfg --exact-size -n 30 --tile-z 16 --device-time
done
C="--x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20"
TAG=prev run $PA $C
TAG=new run $A $C
TAG=new-nozz STENCIL_ASTAROTH_NO_ZIGZAG=1 run $A $C
TAG=new-nokeep STENCIL_NO_L2_KEEP=1 run $A $C
TAG=new-nozz-nokeep STENCIL_NO_L2_KEEP=1 STENCIL_ASTAROTH_NO_ZIGZAG=1 run $A $C
done
M=gpu__time_duration.sum,dram__bytes_read.sum
for v in "prev:$PJ:" "new:$J:" "nozz:$J:STENCIL_JACOBI_NO_ZIGZAG=1" "nokeep:$J:STENCIL_NO_L2_KEEP=1"; do
  tag=${v%%:*}; rest=${v#*:}; exe=${rest%%:*}; envs=${rest#*:}
  env $envs ncu --metrics $M --clock-control none -k regex:jacobi7 -s 4 -c 1 --csv --log-file gpurun_out/ncu_j_$tag.csv $exe 512 512 512 --exact-size -n 8 --tile-z 16 > /dev/null 2>&1
  echo "## ncu jacobi $tag" >> $O; grep jacobi7 gpurun_out/ncu_j_$tag.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
done
for v in "prev:$PA:" "new:$A:" "nozz:$A:STENCIL_ASTAROTH_NO_ZIGZAG=1" "nokeep:$A:STENCIL_NO_L2_KEEP=1"; do
  tag=${v%%:*}; rest=${v#*:}; exe=${rest%%:*}; envs=${rest#*:}
  env $envs ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 1 --csv --log-file gpurun_out/ncu_a_$tag.csv $exe --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
  echo "## ncu astaroth $tag" >> $O; grep astaroth_tma gpurun_out/ncu_a_$tag.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
done
cat $O
