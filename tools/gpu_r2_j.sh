#!/bin/bash
# round 2, session J: grouped ring (one CTA barrier per two row groups) -- tests, isolated timings, C4 / C3 / C5 benches
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python profiles/kernel_bench.py --iters 10 --only na --match "C4 B2048|C3|C5" > gpurun_out/kernel_bench_na_v8.jsonl 2>&1; echo "kbench rc=$?"; cut -c1-230 gpurun_out/kernel_bench_na_v8.jsonl
for c in c4 c3 c5; do timeout 300 python bench.py --no-cpu-baseline --config $c --steps 10 --warmup 3 > gpurun_out/scale_${c}_n1.json 2> gpurun_out/scale_${c}_n1.err; echo "bench $c rc=$?"; cut -c1-330 gpurun_out/scale_${c}_n1.json; tail -2 gpurun_out/scale_${c}_n1.err; done
