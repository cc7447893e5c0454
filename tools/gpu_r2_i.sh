#!/bin/bash
# round 2, session I: 3:2 pooling modes in the fast tails -- full suite, C3-shape kernel bench, C3 / C1 benches
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python profiles/kernel_bench.py --iters 10 --only na --match "C3" > gpurun_out/kernel_bench_na_c3.jsonl 2>&1; echo "kbench rc=$?"; cut -c1-230 gpurun_out/kernel_bench_na_c3.jsonl
for c in c3 c1; do timeout 300 python bench.py --no-cpu-baseline --config $c --steps 20 --warmup 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"; cut -c1-330 gpurun_out/bench_$c.json; tail -2 gpurun_out/bench_$c.err; done
