#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
A=./build/bin/astaroth-sim
run() { echo "## $TAG $*" >> $O; timeout 300 "$@" 2>/dev/null | grep -E "device|astaroth" >> $O; }
timeout 900 ./build/bin/test_cuda "[astaroth]" > "gpurun_out/test_astaroth.log" 2>&1; echo "test rc=$? $(tail -1 gpurun_out/test_astaroth.log)" >> $O
C="--fields 8 --double --star19 --iters 20"
for cfg in "--x 256 --y 256 --z 256" "--x 323 --y 323 --z 162" "--x 406 --y 203 --z 203" "--x 512 --y 256 --z 128"; do
  TAG=cheaplast run $A $cfg $C
  TAG=natural STENCIL_ASTAROTH_NATURAL_ORDER=1 run $A $cfg $C
done
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> $O
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O
cat $O; tail -3 gpurun_out/pytest_gpu.log; python3 -c "
import json
d=json.loads(open('gpurun_out/bench1.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['verified'], d['exchange_us_per_iter'], d['astaroth'], d['e2e']['value'])"
