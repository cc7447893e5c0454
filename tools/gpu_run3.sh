python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for c in c4 c1 c2 c3 c5; do
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"; python -c "
import json,sys
d=json.load(open('gpurun_out/bench_$c.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','final_loss')}, d['e2e']['value'], d['roofline'] and (d['roofline']['kernel'],d['roofline']['achieved'],d['roofline']['frac']), {k:(v.get('GBps'),v['mean_ms']) for k,v in d['kernels'].items()})"; tail -2 gpurun_out/bench_$c.err
done
python tools/profile_step.py c4 1024 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
