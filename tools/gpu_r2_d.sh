#!/bin/bash
# round 2, session D: full GPU suite, smoke, C4 bench + step profile
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
timeout 300 python tools/profile_step.py c4 2048 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
