#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python profiles/kernel_bench.py --only na > gpurun_out/kernel_bench_na.jsonl 2> gpurun_out/kernel_bench_na.err; echo "kb rc=$?"
for c in c4 c3; do
timeout 600 python bench.py --config $c --no-cpu-baseline > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"
done
