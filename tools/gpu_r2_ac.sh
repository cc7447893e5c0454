#!/bin/bash
# round 2, session AC: lane-parallel general skip gradient in the generic row-wise backward (C2 widths 256 -> 48)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py tests/test_gpu_amp.py -q -m gpu -x 2>&1 | grep -E "Error|FAILED|passed|failed" | head
timeout 300 python profiles/kernel_bench.py --iters 10 --only na --match "C2|C3 GELU" > gpurun_out/kb_ac.jsonl 2>&1; python - <<'PY'
import json
for l in open("gpurun_out/kb_ac.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print("  %-14s %-44s %8.4f ms  %.3f" % (d["kernel"], d["shape"], d["ms_median"], d["frac_of_measured_peak"]))
PY
for c in c2 c1; do timeout 300 python bench.py --no-cpu-baseline --config $c --steps 20 --warmup 5 > gpurun_out/bench_${c}_v13.json 2> gpurun_out/bench_${c}_v13.err; echo "bench $c rc=$?"; cut -c1-250 gpurun_out/bench_${c}_v13.json; done
