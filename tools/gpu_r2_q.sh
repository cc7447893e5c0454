#!/bin/bash
# round 2, session Q: ring kernels with the new defaults; group-size and packed-math A/B; ncu --set full of the ring kernels
set -u
mkdir -p gpurun_out /tmp/ncu
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py tests/test_zz_no_alloc.py -q -m gpu -k "fftmix or noalloc or guard" -x 2>&1 | tail -5
M="f32 B1024|C2 7x7|C1 8x8"
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only fft --match "$M" > gpurun_out/kb_q_$tag.jsonl 2> gpurun_out/kb_q_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_q_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-32s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb default A=1
kb g2 SB200_FFT_RING_G=2
kb g4 SB200_FFT_RING_G=4
kb fwd8_bwd8 SB200_FFT_RING_G=8
kb packed SB200_LIB_NAME=libspectre_b200_pk.so
kb packed_g2 SB200_LIB_NAME=libspectre_b200_pk.so SB200_FFT_RING_G=2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fftmix_ring" -c 8 -o /tmp/ncu/ring -f \
  python profiles/kernel_bench.py --iters 1 --only fft --match "f32 B1024|C2 7x7" > gpurun_out/ncu_ring.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/ring.ncu-rep --page raw --csv > gpurun_out/ncu_ring_raw.csv 2>/dev/null
gzip -f gpurun_out/ncu_ring_raw.csv
