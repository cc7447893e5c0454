#!/bin/bash
# tests + norm_act kernel bench + benches + full-set ncu capture of the mixers at the bench batch (B=2048)
set -u
mkdir -p gpurun_out /tmp/ncu
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python profiles/kernel_bench.py --only na > gpurun_out/kernel_bench_na.jsonl 2> gpurun_out/kernel_bench_na.err; echo "kb rc=$?"
for c in c4 c3 c1 c2; do
  timeout 600 python bench.py --config $c --no-cpu-baseline > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_tc" -c 8 -o /tmp/ncu/mix2048 -f \
  python profiles/kernel_bench.py --iters 1 --only fftdwt --match "B2048" > gpurun_out/ncu_full_2048.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/mix2048.ncu-rep --page raw --csv 2>/dev/null | gzip > gpurun_out/ncu_mixers_b2048_raw.csv.gz
du -sh gpurun_out
