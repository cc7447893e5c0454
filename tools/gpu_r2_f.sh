#!/bin/bash
# round 2, session F: new tests (guards, FNet / FFT module), C5 with the reference's DINOv3 teacher, C1 / C2 benches
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline --config c5 --steps 10 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"; cut -c1-700 gpurun_out/bench_c5.json; tail -5 gpurun_out/bench_c5.err
for c in c1 c2; do timeout 600 python bench.py --no-cpu-baseline --config $c --steps 10 --warmup 3 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"; cut -c1-330 gpurun_out/bench_$c.json; tail -2 gpurun_out/bench_$c.err; done
