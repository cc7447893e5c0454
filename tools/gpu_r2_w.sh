#!/bin/bash
# round 2, session W: multi-level Haar register kernel, bf16: ring depth and TMA store vs per-thread global stores
set -u
mkdir -p gpurun_out
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only dwtml --match "bf16" > gpurun_out/kb_w_$tag.jsonl 2> gpurun_out/kb_w_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_w_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-36s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb s3 A=1
kb s2 SB200_DWTML_S=2
kb s4 SB200_DWTML_S=4
kb stg_s3 SB200_DWTML_STG=1
kb stg_s2 SB200_DWTML_STG=1 SB200_DWTML_S=2
SB200_DWTML_STG=1 timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "multilevel" 2>&1 | tail -2
