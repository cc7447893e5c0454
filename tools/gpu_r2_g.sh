#!/bin/bash
# round 2, session G: CUDA-graph step: test + C3 / C1 / C2 / C4 benches with and without the graph
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py -m gpu -q -x -k "graph or patch_embed or norm_act_dropout" > gpurun_out/pytest_graph.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_graph.log
for c in c3 c1 c2; do
  timeout 600 python bench.py --no-cpu-baseline --config $c --steps 20 --warmup 5 --cuda-graph > gpurun_out/bench_${c}_graph.json 2> gpurun_out/bench_${c}_graph.err; echo "bench $c graph rc=$?"; cut -c1-330 gpurun_out/bench_${c}_graph.json; tail -3 gpurun_out/bench_${c}_graph.err
done
timeout 600 python bench.py --no-cpu-baseline --cuda-graph > gpurun_out/bench_c4_graph.json 2> gpurun_out/bench_c4_graph.err; echo "bench c4 graph rc=$?"; cut -c1-330 gpurun_out/bench_c4_graph.json; tail -3 gpurun_out/bench_c4_graph.err
timeout 600 python bench.py --no-cpu-baseline --batch 256 --cuda-graph > gpurun_out/bench_c4_b256_graph.json 2>> gpurun_out/bench_c4_graph.err; echo "bench c4 b256 graph rc=$?"; cut -c1-330 gpurun_out/bench_c4_b256_graph.json
