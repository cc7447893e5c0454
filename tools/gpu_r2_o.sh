#!/bin/bash
# round 2, session O: TMA-fed in-place fp32 FFT mixers (fftmix_ring.cu) -- parity, then A/B against the first-generation kernels
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py tests/test_zz_no_alloc.py -q -m gpu -k "fftmix or noalloc or guard" -x 2>&1 | tail -15
M="f32 B1024|C2 7x7|C1 8x8"
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only fft --match "$M" > gpurun_out/kb_ring_$tag.jsonl 2> gpurun_out/kb_ring_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_ring_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-32s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb old SB200_FFT_RING=0
kb ring SB200_FFT_RING=1
kb ring_s2 SB200_FFT_RING_STAGES=2
kb ring_g4 SB200_FFT_RING_G=4
kb ring_g4_s2 SB200_FFT_RING_G=4 SB200_FFT_RING_STAGES=2
kb ring_g8 SB200_FFT_RING_G=8
kb old_nopre SB200_KB_PREROLL=0 SB200_FFT_RING=0
kb ring_nopre SB200_KB_PREROLL=0
