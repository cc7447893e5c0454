#!/bin/bash
# usage: tools/gpu_scale2.sh N cfg [cfg ...]   -- bench.py under torchrun at N ranks (the way the driver launches it), one config after
# the other; every run under its own short timeout (a hang must not burn N x the box time)
set -u
N=$1; shift
mkdir -p gpurun_out
port=29531
for c in "$@"; do
  port=$((port + 1))
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --config $c --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/scale_${c}_n1.json 2> gpurun_out/scale_${c}_n1.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --config $c --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_${c}_n$N.json 2> gpurun_out/scale_${c}_n$N.err
  fi
  echo "$c n$N rc=$?"
  grep '"metric"' gpurun_out/scale_${c}_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['name'], d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])" || tail -5 gpurun_out/scale_${c}_n$N.err
done
