#!/bin/bash
# round 2: ncu evidence of the final kernels -- (1) launch list of the bench command (eager, so kernel names map 1:1),
# (2) --set full of the mixers and the row-wise tails at the bench batch (raw CSV exported on the box).
set -u
mkdir -p gpurun_out /tmp/ncu
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/ncu_launches_c4.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cuda-graph > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
gzip -f gpurun_out/ncu_launches_c4.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fftmix_tc|dwtmix" -c 12 -o /tmp/ncu/mixers -f \
  python profiles/kernel_bench.py --iters 1 --only fftdwt --match "B2048|C3 dwt" > gpurun_out/ncu_full_mix.log 2>&1; echo "ncu mixers rc=$?"
ncu -i /tmp/ncu/mixers.ncu-rep --page raw --csv > gpurun_out/ncu_mixers_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fast_fwd|fast_bwd|patch_gather|embed_finish" -c 24 -o /tmp/ncu/na -f \
  python profiles/kernel_bench.py --iters 1 --only na --match "C4 B2048" > gpurun_out/ncu_full_na.log 2>&1; echo "ncu na rc=$?"
ncu -i /tmp/ncu/na.ncu-rep --page raw --csv > gpurun_out/ncu_na_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_mixers_raw.csv gpurun_out/ncu_na_raw.csv
du -sh gpurun_out
