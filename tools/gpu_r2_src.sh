#!/bin/bash
# round 2: SASS-level stall samples (ncu source page) of the kernels that sit below their roofline for no stated reason
set -u
mkdir -p gpurun_out /tmp/ncu
cap() { tag=$1; kre=$2; only=$3; match=$4; skip=$5
  timeout 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$kre" -c 1 -s $skip -o /tmp/ncu/$tag -f \
    python profiles/kernel_bench.py --iters 1 --only $only --match "$match" > gpurun_out/ncu_src_$tag.log 2>&1; echo "ncu $tag rc=$?"
  ncu -i /tmp/ncu/$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_src_$tag.csv 2>/dev/null
  ncu -i /tmp/ncu/$tag.ncu-rep --page raw --csv > gpurun_out/ncu_raw_$tag.csv 2>/dev/null
  gzip -f gpurun_out/ncu_src_$tag.csv gpurun_out/ncu_raw_$tag.csv
  python tools/ncu_src_top.py gpurun_out/ncu_src_$tag.csv.gz 10
}
cap tcfwd256 "fftmix_tc_fwd_kernel" fft "C4 14x14 E384 bf16 B256" 3
cap tcbwd256 "fftmix_tc_bwd_kernel" fft "C4 14x14 E384 bf16 B256" 3
cap dwtfwd "dwtmix.*kernel" dwt "C3 dwt 8x8" 3
cap dwtbwd "dwtmix.*kernel" dwt "C3 dwt 8x8" 7
cap ringc2 "fftmix_ring_fwd" fft "C2 7x7" 3
cap gelufwd "fast_fwd_kernel<3, 64" na "C4 B2048 GELU.LN.h1" 3
cap gelubwd "fast_bwd_kernel" na "C4 B2048 GELU.LN.h1" 3
