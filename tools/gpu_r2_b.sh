#!/bin/bash
# round 2, session B: fast row-wise tails -- parity, isolated timings (fast vs generic), C4 bench
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_bench_shapes.py tests/test_gpu_amp.py -q -x -k "norm_act or encoder or stacked or amp" > gpurun_out/pytest_na.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_na.log
timeout 600 python profiles/kernel_bench.py --iters 10 --only na --match "C4 B2048|C5|C4 B256 LN1|C4 B256 GELU" > gpurun_out/kernel_bench_na_fast.jsonl 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench_na_fast.jsonl
SB200_NA_FAST=0 timeout 600 python profiles/kernel_bench.py --iters 10 --only na --match "C4 B2048" > gpurun_out/kernel_bench_na_generic.jsonl 2>&1; echo "kbench generic rc=$?"; cat gpurun_out/kernel_bench_na_generic.jsonl
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
timeout 300 python tools/profile_step.py c4 2048 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
