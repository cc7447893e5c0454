python bench.py --steps 10 --warmup 3 --batch 2048 --no-cpu-baseline > gpurun_out/bench_c4_b2048.json 2> gpurun_out/bench_c4_b2048.err; echo "rc=$?"; python -c "
import json
d=json.load(open('gpurun_out/bench_c4_b2048.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','final_loss')}, d['e2e'], d['roofline'], d['kernels'])"; tail -2 gpurun_out/bench_c4_b2048.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
python - <<'PY'
import torch,sys
sys.path.insert(0,'.')
from vit_spectre_experiments_b200.configs import load_named
from vit_spectre_experiments_b200.train import Trainer
cfg=load_named('c4'); tr=Trainer(cfg,'cuda:0'); img,lab=tr.synthetic_batch(2048)
for _ in range(2): tr.step(img,lab)
torch.cuda.synchronize(); print('peak mem GB', torch.cuda.max_memory_allocated()/1e9)
PY
