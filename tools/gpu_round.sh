#!/bin/bash
# GPU-box session 1: tests -> benches (all configs) -> step profile -> kernel bench -> ncu launch list.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_c4_reference.json 2>> gpurun_out/bench_c4.err; echo "bench ref rc=$?"; cat gpurun_out/bench_c4_reference.json
for b in 256 1024; do python bench.py --steps 20 --warmup 5 --batch $b --no-cpu-baseline > gpurun_out/bench_c4_b$b.json 2>> gpurun_out/bench_c4.err; echo "bench$b rc=$?"; cut -c1-200 gpurun_out/bench_c4_b$b.json; done
for c in c1 c2 c3 c5; do python bench.py --steps 10 --warmup 3 --config $c > gpurun_out/bench_$c.json 2>> gpurun_out/bench_c4.err; echo "bench $c rc=$?"; cut -c1-200 gpurun_out/bench_$c.json; done
python tools/profile_step.py c4 2048 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
python profiles/kernel_bench.py --iters 20 > gpurun_out/kernel_bench.jsonl 2>&1; echo "kbench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/ncu_launches_c4.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
gzip -f gpurun_out/ncu_launches_c4.csv
du -sh gpurun_out
