#!/bin/bash
# Scratch script for one-off GPU experiments (edited per experiment, run through gpurun).
mkdir -p gpurun_out
cd /root/repo
export STENCIL_LOG_LEVEL=1
O=gpurun_out/exp.txt
: > $O
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/pytest_gpu.log)" >> $O
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" >> $O 2>&1
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> $O
timeout 900 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?" >> $O
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default-args rc=$?" >> $O
cat $O; python3 -c "
import json
for f in ['bench1','bench_ref','bench_default']:
    d=json.loads(open('gpurun_out/'+f+'.json').read().strip().splitlines()[-1])
    print(f, d['value'], d['ms_per_step'], d.get('verified'), d['exchange_us_per_iter'], (d.get('astaroth') or {}).get('ms_per_step_device'), d['e2e']['value'], d['clocks'])"
