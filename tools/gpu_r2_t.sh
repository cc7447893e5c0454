#!/bin/bash
# round 2, session T: register-resident multi-level Haar (dwtmix_ml_reg.cu): parity, A/B against dwtmix_ml.cu
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_fused.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | grep -E "Error|FAILED|passed|failed" | head
kb() { tag=$1; shift; env "$@" timeout 200 python profiles/kernel_bench.py --iters 30 --only dwtml > gpurun_out/kb_t_$tag.jsonl 2> gpurun_out/kb_t_$tag.err; echo "== $tag rc=$?"; python - <<PY
import json
for l in open("gpurun_out/kb_t_$tag.jsonl"):
    d = json.loads(l); print("  %-12s %-36s %8.4f ms  %6.1f GB/s  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["GBps_median"], d["frac_of_measured_peak"]))
PY
}
kb old SB200_DWTML_REG=0
kb reg A=1
