#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
A=./build/bin/astaroth-sim
OLD=./build/old_src/bld/bin/astaroth-sim
run() { echo "## $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth" >> $O; }
for i in 1 2; do
run $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
run $OLD --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 20
done
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active
J="./build/bin/jacobi3d"
for cfg in "512 512 512" "1024 512 256" "2048 256 256"; do
  for mode in "" "--no-overlap"; do
    tag=$(echo "$cfg$mode" | tr ' ' '_')
    ncu --metrics $M --clock-control none -k regex:jacobi7 -s 4 -c 3 --csv --log-file gpurun_out/ncu_$tag.csv $J $cfg --exact-size -n 8 --tile-z 16 $mode > /dev/null 2>&1
    echo "## ncu $cfg $mode" >> $O; grep -E "jacobi7" gpurun_out/ncu_$tag.csv | awk -F'","' '{print $5, $(NF-2), $(NF)}' | tail -15 >> $O
  done
done
ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 2 --csv --log-file gpurun_out/ncu_ast_new.csv $A --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
ncu --metrics $M --clock-control none -k regex:astaroth_tma -s 3 -c 2 --csv --log-file gpurun_out/ncu_ast_old.csv $OLD --x 256 --y 256 --z 256 --fields 8 --double --star19 --iters 6 > /dev/null 2>&1
echo "## ncu astaroth new" >> $O; grep astaroth_tma gpurun_out/ncu_ast_new.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
echo "## ncu astaroth old" >> $O; grep astaroth_tma gpurun_out/ncu_ast_old.csv | awk -F'","' '{print $(NF-2), $(NF)}' >> $O
cat $O
