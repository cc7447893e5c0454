#!/bin/bash
# round 2, session H (2 GPUs): graph-mode step under torchrun (in-graph NCCL all-reduce) vs eager DDP, C4 and C3
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fused.py -m gpu -q -x -k "graph" > gpurun_out/pytest_graph.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_graph.log
for c in c4 c3; do
  for mode in "" "--no-cuda-graph"; do
    tag=${mode:+_eager}
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --config $c --gpus 2 --steps 10 --warmup 3 $mode > gpurun_out/scale_${c}_n2${tag}.json 2> gpurun_out/scale_${c}_n2${tag}.err; echo "$c n2 $mode rc=$?"
    cut -c1-330 gpurun_out/scale_${c}_n2${tag}.json; tail -3 gpurun_out/scale_${c}_n2${tag}.err
  done
done
