#!/bin/bash
# round 2, session X: where do the warps of the register-resident Haar kernel stall? (ncu source page, SASS level)
set -u
mkdir -p gpurun_out /tmp/ncu
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"dwtml_reg" -c 1 -s 3 -o /tmp/ncu/mlsrc -f \
  python profiles/kernel_bench.py --iters 1 --only dwtml --match "bf16" > gpurun_out/ncu_mlsrc.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/mlsrc.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_mlsrc_sass.csv 2>gpurun_out/ncu_mlsrc_sass.err; echo "export rc=$?"
gzip -f gpurun_out/ncu_mlsrc_sass.csv; ls -la gpurun_out/ncu_mlsrc_sass.csv.gz
