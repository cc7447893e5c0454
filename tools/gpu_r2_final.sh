#!/bin/bash
# final check of a build: full GPU suite, smoke, default bench (+ reference arm)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log | cut -c1-250
timeout 600 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_c4.json; tail -2 gpurun_out/bench_c4.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_c4_reference.json 2> gpurun_out/bench_c4_reference.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_c4_reference.json; tail -2 gpurun_out/bench_c4_reference.err
