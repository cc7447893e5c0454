#!/bin/bash
# round 2, session L: asynchronous ring hand-back (default build) vs the grouped CTA barrier (libspectre_b200_sync.so)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py -m gpu -q > gpurun_out/pytest_async.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_async.log | cut -c1-200
timeout 300 python profiles/kernel_bench.py --iters 15 --only na --match "C4 B2048|C5|C3" > gpurun_out/kb_async.jsonl 2>&1; echo "kb async rc=$?"
SB200_LIB_NAME=libspectre_b200_sync.so timeout 300 python profiles/kernel_bench.py --iters 15 --only na --match "C4 B2048|C5|C3" > gpurun_out/kb_sync.jsonl 2>&1; echo "kb sync rc=$?"
python - <<'PY'
import json
a=[json.loads(l) for l in open('gpurun_out/kb_async.jsonl') if l.startswith('{')]
b=[json.loads(l) for l in open('gpurun_out/kb_sync.jsonl') if l.startswith('{')]
for x,y in zip(a,b):
    print(f"{x['kernel'][-3:]} {x['shape']:46s} async {x['ms_median']:.4f} ms ({x['frac_of_measured_peak']:.3f})   sync {y['ms_median']:.4f} ms ({y['frac_of_measured_peak']:.3f})")
PY
timeout 300 python bench.py --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/bench_c4_async.json 2> gpurun_out/bench_c4_async.err; echo "bench async rc=$?"; cut -c1-200 gpurun_out/bench_c4_async.json
SB200_LIB_NAME=libspectre_b200_sync.so timeout 300 python bench.py --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/bench_c4_sync.json 2> gpurun_out/bench_c4_sync.err; echo "bench sync rc=$?"; cut -c1-200 gpurun_out/bench_c4_sync.json
