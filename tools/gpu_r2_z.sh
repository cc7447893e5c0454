#!/bin/bash
# round 2, session Z: register Haar kernel with direct stores as the default (all dtypes / grids): parity + A/B
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py -q -m gpu -x -k "multilevel or haar or guard" 2>&1 | grep -E "Error|FAILED|passed|failed" | head
SB200_DWTML_STG=0 timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "multilevel" 2>&1 | grep -E "Error|FAILED|passed|failed" | head
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only dwtml > gpurun_out/kb_z_$tag.jsonl 2> gpurun_out/kb_z_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_z_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-36s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb direct A=1
kb tma SB200_DWTML_STG=0
