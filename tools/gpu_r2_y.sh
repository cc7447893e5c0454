#!/bin/bash
# round 2, session Y: batched prologue loads (weight tile of the register Haar kernel, filter tile of the ring FFT kernels)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -E "Error|FAILED|passed|failed" | head
kb() { tag=$1; only=$2; match=$3; shift 3; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only $only --match "$match" > gpurun_out/kb_y_$tag.jsonl 2> gpurun_out/kb_y_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_y_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-36s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb fft fft "f32 B1024|C2 7x7|f32 B256|C1 8x8" A=1
kb ml dwtml "dwtml" A=1
kb ml_stg dwtml "bf16" SB200_DWTML_STG=1
kb ml_s2 dwtml "bf16" SB200_DWTML_S=2
