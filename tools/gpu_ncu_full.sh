#!/bin/bash
# GPU-box session 2: ncu --set full captures of every kernel; exported to CSV on the box (the .ncu-rep files are too
# large to travel back), raw page for all launches + source page for the two tensor-core mixers.
set -u
mkdir -p gpurun_out /tmp/ncu
ncu --set full --clock-control none --import-source on -k regex:"fftmix_tc|dwtmix|fftmix_(fwd|bwd)_kernel" -c 48 -o /tmp/ncu/mixers -f \
  python profiles/kernel_bench.py --iters 1 --only fftdwt --match "B1024|B2048|B4096" > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/ncu/mixers.ncu-rep --page raw --csv > gpurun_out/ncu_mixers_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:"norm_act|reduce_" -c 40 -o /tmp/ncu/normact -f \
  python profiles/kernel_bench.py --iters 1 --only na --match "C4" > gpurun_out/ncu_full_na.log 2>&1; echo "ncu na rc=$?"
ncu -i /tmp/ncu/normact.ncu-rep --page raw --csv > gpurun_out/ncu_normact_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_mixers_raw.csv gpurun_out/ncu_normact_raw.csv
du -sh gpurun_out
