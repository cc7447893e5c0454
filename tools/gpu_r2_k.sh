#!/bin/bash
# round 2, session K: A/B of the wide backward's thread shape (2 vs 3 vectors per thread)
set -u
mkdir -p gpurun_out
SB200_NA_BWD_VPT3=1 timeout 600 python -m pytest tests/test_gpu_fused.py tests/test_gpu_bench_shapes.py tests/test_gpu_guards.py -m gpu -q -k "norm_act or encoder" > gpurun_out/pytest_vpt3.log 2>&1; echo "pytest vpt3 rc=$?"; tail -3 gpurun_out/pytest_vpt3.log | cut -c1-200
timeout 300 python profiles/kernel_bench.py --iters 15 --only na --match "B2048 GELU\(LN\(h1|C5 B512 GELU\(LN\(h1" > gpurun_out/kb_vpt2.jsonl 2>&1; cut -c1-230 gpurun_out/kb_vpt2.jsonl
SB200_NA_BWD_VPT3=1 timeout 300 python profiles/kernel_bench.py --iters 15 --only na --match "B2048 GELU\(LN\(h1|C5 B512 GELU\(LN\(h1" > gpurun_out/kb_vpt3.jsonl 2>&1; cut -c1-230 gpurun_out/kb_vpt3.jsonl
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
