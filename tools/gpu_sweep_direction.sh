#!/bin/bash
set -u
mkdir -p gpurun_out
for rev in 0000 1100 0011 0010 0001 0000; do
timeout 300 python bench.py --no-cpu-baseline --steps 12 --warmup 4 > gpurun_out/bench_rev_$rev.json 2> gpurun_out/bench_rev.err; echo "rev $rev rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bench_rev_$rev.json')); print('$rev', d['value'], d['ms_per_step'], {k:v['mean_ms'] for k,v in d['kernels'].items()})"
done
