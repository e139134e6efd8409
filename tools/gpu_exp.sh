#!/bin/bash
# scratch: Astaroth thin z chunks at the end of the table, one GPU
mkdir -p gpurun_out
A="./build/bin/astaroth-sim --fields 8 --double --star19 --iters 20"
{
  timeout 60 ./build/bin/test_cuda "[tail]" 2>&1 | tail -3
  for sz in "256 256 256" "323 323 162" "512 256 128"; do
    set -- $sz
    for c in "0 8" "16 8" "32 8" "24 12" "48 8"; do
      r=$(STENCIL_ASTAROTH_TAIL_PLANES=${c% *} STENCIL_ASTAROTH_TAIL_DEPTH=${c#* } timeout 30 $A --x $1 --y $2 --z $3 2>&1 | grep "^astaroth-sim" | awk -F, '{print $12*1e3}')
      echo "$sz tail $c : $r"
    done
  done
} > gpurun_out/exp.txt 2>&1
cat gpurun_out/exp.txt
