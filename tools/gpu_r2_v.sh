#!/bin/bash
# round 2, session V: C1 / C2 / C3 benches with the ring FFT kernels; ncu --set full of the register-resident multi-level Haar
set -u
mkdir -p gpurun_out /tmp/ncu
for c in c1 c2 c3; do timeout 300 python bench.py --no-cpu-baseline --config $c --steps 20 --warmup 5 > gpurun_out/bench_${c}_v11.json 2> gpurun_out/bench_${c}_v11.err; echo "bench $c rc=$?"; cut -c1-330 gpurun_out/bench_${c}_v11.json; tail -2 gpurun_out/bench_${c}_v11.err; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dwtml_reg" -c 16 -o /tmp/ncu/mlreg -f \
  python profiles/kernel_bench.py --iters 1 --only dwtml > gpurun_out/ncu_mlreg.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/mlreg.ncu-rep --page raw --csv > gpurun_out/ncu_mlreg_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_mlreg_raw.csv
