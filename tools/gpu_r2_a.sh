#!/bin/bash
# round 2, session A: full GPU suite (incl. the new bench-shape / AMP tests) + row-kernel baseline timings
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python profiles/kernel_bench.py --iters 10 --only na > gpurun_out/kernel_bench_na.jsonl 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench_na.jsonl
./profiles/microbench/pipes2 > gpurun_out/pipes2.txt 2>&1; echo "pipes2 rc=$?"; cat gpurun_out/pipes2.txt
