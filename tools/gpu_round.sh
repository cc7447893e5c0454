#!/bin/bash
# One GPU-box session: tests -> bench -> step profile -> kernel bench -> ncu launch list -> ncu full captures.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json
python bench.py --steps 20 --warmup 5 --batch 256 --no-cpu-baseline > gpurun_out/bench_c4_b256.json 2>> gpurun_out/bench_c4.err; echo "bench256 rc=$?"; cut -c1-200 gpurun_out/bench_c4_b256.json
python bench.py --steps 20 --warmup 5 --config c3 --no-cpu-baseline > gpurun_out/bench_c3.json 2>> gpurun_out/bench_c4.err; echo "bench c3 rc=$?"; cut -c1-200 gpurun_out/bench_c3.json
python tools/profile_step.py c4 1024 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
python profiles/kernel_bench.py --iters 20 > gpurun_out/kernel_bench.jsonl 2>&1; echo "kbench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/ncu_launches_c4.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"fftmix_tc|dwtmix|fftmix_(fwd|bwd)_kernel" -c 40 -o gpurun_out/ncu_mixers -f \
  python profiles/kernel_bench.py --iters 1 --match "B1024" > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
