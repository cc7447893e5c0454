#!/bin/bash
# SASS-level stall samples of the generic fp32 tails at the C2 widths (48 <-> 256, table-driven pooling)
set -u
mkdir -p gpurun_out /tmp/ncu
cap() { tag=$1; kre=$2; match=$3; skip=$4
  timeout 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$kre" -c 1 -s $skip -o /tmp/ncu/$tag -f \
    python profiles/kernel_bench.py --iters 1 --only na --match "$match" > gpurun_out/ncu_src_$tag.log 2>&1; echo "ncu $tag rc=$?"
  ncu -i /tmp/ncu/$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_src_$tag.csv 2>/dev/null
  ncu -i /tmp/ncu/$tag.ncu-rep --page raw --csv > gpurun_out/ncu_raw_$tag.csv 2>/dev/null
  gzip -f gpurun_out/ncu_src_$tag.csv gpurun_out/ncu_raw_$tag.csv
  python tools/ncu_src_top.py gpurun_out/ncu_src_$tag.csv.gz 14
  python tools/ncu_summarise.py gpurun_out/ncu_raw_$tag.csv.gz | cut -c1-300
}
cap c2bwd256 "norm_act_bwd_kernel" "C2 GELU.LN.h1" 3
cap c2fwd256 "norm_act_fwd_kernel" "C2 GELU.LN.h1" 3
cap c2bwd48 "norm_act_bwd_kernel" "C2 GELU.LN.h2" 3
