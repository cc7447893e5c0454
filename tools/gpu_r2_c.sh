#!/bin/bash
# round 2, session C: ncu --set full of the fast row-wise tails at the bench batch (raw CSV + source page of two kernels)
set -u
mkdir -p gpurun_out /tmp/ncu
ncu --set full --clock-control none --import-source on -k regex:"fast_fwd|fast_bwd|norm_act" -c 24 -o /tmp/ncu/na -f \
  python profiles/kernel_bench.py --iters 1 --only na --match "C4 B2048" > gpurun_out/ncu_na.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_na.log
ncu -i /tmp/ncu/na.ncu-rep --page raw --csv > gpurun_out/ncu_na_raw.csv 2>/dev/null
ncu -i /tmp/ncu/na.ncu-rep --page source --csv -k regex:"fast_fwd_kernel<3, 64" > gpurun_out/ncu_na_src_fwd1536.csv 2>/dev/null
ncu -i /tmp/ncu/na.ncu-rep --page details -k regex:"fast_fwd_kernel<3, 64" 2>/dev/null | head -400 > gpurun_out/ncu_na_details_fwd1536.txt
ncu -i /tmp/ncu/na.ncu-rep --page details -k regex:"fast_fwd_kernel<3, 16" 2>/dev/null | head -400 > gpurun_out/ncu_na_details_fwd384.txt
gzip -f gpurun_out/ncu_na_raw.csv gpurun_out/ncu_na_src_fwd1536.csv
du -sh gpurun_out
