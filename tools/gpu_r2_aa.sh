#!/bin/bash
# round 2, session AA: single-level Haar backward on 4x4 / 8x8 grids through the register kernel (sub-band weight layout)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | grep -E "Error|FAILED|passed|failed" | head
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only dwt --match "8x8" > gpurun_out/kb_aa_$tag.jsonl 2> gpurun_out/kb_aa_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_aa_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-36s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb old SB200_DWT_REG_BWD=0
kb reg A=1
timeout 300 python bench.py --no-cpu-baseline --config c3 --steps 20 --warmup 5 > gpurun_out/bench_c3_v12.json 2> gpurun_out/bench_c3_v12.err; echo "bench c3 rc=$?"; cut -c1-260 gpurun_out/bench_c3_v12.json
