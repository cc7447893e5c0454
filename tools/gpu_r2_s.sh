#!/bin/bash
# round 2, session S: generalized ring kernels (precomputed indices, one-channel lanes for 14x14 / 16x16): parity, A/B, ncu
set -u
mkdir -p gpurun_out /tmp/ncu
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py tests/test_gpu_amp.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -5
M="f32 B1024|C2 7x7|C1 8x8|f32 B256"
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only fft --match "$M" > gpurun_out/kb_s_$tag.jsonl 2> gpurun_out/kb_s_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_s_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-32s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb old SB200_FFT_RING=0
kb ring A=1
kb ring_g2 SB200_FFT_RING_G=2
kb ring_s3 SB200_FFT_RING_STAGES=3
SB200_FFT_RING=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_fwd_kernel|fftmix_bwd_kernel" -c 24 -o /tmp/ncu/old -f \
  python profiles/kernel_bench.py --iters 1 --only fft --match "f32 B1024|C2 7x7|f32 B256" > gpurun_out/ncu_old.log 2>&1; echo "ncu old rc=$?"
ncu -i /tmp/ncu/old.ncu-rep --page raw --csv > gpurun_out/ncu_fft32_old_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_ring" -c 24 -o /tmp/ncu/ring -f \
  python profiles/kernel_bench.py --iters 1 --only fft --match "f32 B1024|C2 7x7|f32 B256" > gpurun_out/ncu_ring.log 2>&1; echo "ncu ring rc=$?"
ncu -i /tmp/ncu/ring.ncu-rep --page raw --csv > gpurun_out/ncu_fft32_ring_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_fft32_old_raw.csv gpurun_out/ncu_fft32_ring_raw.csv
