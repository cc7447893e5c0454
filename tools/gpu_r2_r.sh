#!/bin/bash
# round 2, session R: C2 (7x7, E=48) as one 48-channel tile (P=24) against three 16-channel tiles; ncu of the C2 ring kernels
set -u
mkdir -p gpurun_out /tmp/ncu
M="C2 7x7"
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 40 --only fft --match "$M" > gpurun_out/kb_r_$tag.jsonl 2> gpurun_out/kb_r_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_r_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-32s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb p8_g4 A=1
kb p24_g2 SB200_FFT_RING_P=24 SB200_FFT_RING_G=2
kb p24_g4 SB200_FFT_RING_P=24 SB200_FFT_RING_G=4
kb p24_g4_s3 SB200_FFT_RING_P=24 SB200_FFT_RING_G=4 SB200_FFT_RING_STAGES=3
SB200_FFT_RING_P=24 SB200_FFT_RING_G=2 timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shapes.py -q -m gpu -k "fftmix" -x 2>&1 | tail -3
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_ring" -c 4 -o /tmp/ncu/ringc2 -f \
  python profiles/kernel_bench.py --iters 1 --only fft --match "C2 7x7" > gpurun_out/ncu_ringc2.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/ringc2.ncu-rep --page raw --csv > gpurun_out/ncu_ringc2_raw.csv 2>/dev/null
SB200_FFT_RING_P=24 SB200_FFT_RING_G=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_ring" -c 4 -o /tmp/ncu/ringc2p24 -f \
  python profiles/kernel_bench.py --iters 1 --only fft --match "C2 7x7" > gpurun_out/ncu_ringc2p24.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/ringc2p24.ncu-rep --page raw --csv > gpurun_out/ncu_ringc2p24_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_ringc2_raw.csv gpurun_out/ncu_ringc2p24_raw.csv
