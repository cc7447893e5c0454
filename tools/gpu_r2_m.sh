#!/bin/bash
# round 2, session M: sub-warp rows in the generic row-wise kernels (N = 48 of config C2): full suite, C2-shape kernel bench, C2 / C1 benches
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python profiles/kernel_bench.py --iters 10 --only na --match "C2" > gpurun_out/kb_c2.jsonl 2>&1; cut -c1-230 gpurun_out/kb_c2.jsonl
for c in c2 c1; do timeout 300 python bench.py --no-cpu-baseline --config $c --steps 20 --warmup 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"; cut -c1-330 gpurun_out/bench_$c.json; tail -2 gpurun_out/bench_$c.err; done
