#!/bin/bash
# scaling run on one box: bench.py launched the way the driver does, for every N given
for N in "$@"; do
if [ "$N" = "1" ]; then
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 rc=$?"
else
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "n$N rc=$?"
fi
grep '"metric"' gpurun_out/scale_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'])"
done
# the DWT config (C3) at the largest N given, for the record
if [ "${N:-1}" != "1" ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --config c3 --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_c3_n$N.json 2> gpurun_out/scale_c3_n$N.err; echo "c3 n$N rc=$?"
grep '"metric"' gpurun_out/scale_c3_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3', d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'])"
fi
