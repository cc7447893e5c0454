#!/bin/bash
# round 2, session E: PDL launches + strip-staged DWT: full suite, mixer kernel bench, C4 / C3 bench (+ PDL off for A/B)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python profiles/kernel_bench.py --iters 10 --only fftdwt > gpurun_out/kernel_bench_mix.jsonl 2>&1; echo "kbench rc=$?"; cut -c1-230 gpurun_out/kernel_bench_mix.jsonl
SB200_DWT_STRIP=0 timeout 300 python profiles/kernel_bench.py --iters 10 --only dwt --match "C3|f32" > gpurun_out/kernel_bench_dwt_nostrip.jsonl 2>&1; cut -c1-230 gpurun_out/kernel_bench_dwt_nostrip.jsonl
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
SB200_PDL=0 timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c4_nopdl.json 2>> gpurun_out/bench_c4.err; echo "bench nopdl rc=$?"; cut -c1-330 gpurun_out/bench_c4_nopdl.json
timeout 600 python bench.py --no-cpu-baseline --batch 256 > gpurun_out/bench_c4_b256.json 2>> gpurun_out/bench_c4.err; echo "bench b256 rc=$?"; cut -c1-330 gpurun_out/bench_c4_b256.json
timeout 600 python bench.py --no-cpu-baseline --config c3 > gpurun_out/bench_c3.json 2>> gpurun_out/bench_c4.err; echo "bench c3 rc=$?"; cut -c1-330 gpurun_out/bench_c3.json
