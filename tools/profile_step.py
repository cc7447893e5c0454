"""torch.profiler table of one training step (which kernels dominate outside the mixers).
Usage: python tools/profile_step.py [c4] [batch]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vit_spectre_experiments_b200.configs import load_named
from vit_spectre_experiments_b200.train import Trainer

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
cfg = load_named(name)
batch = int(sys.argv[2]) if len(sys.argv) > 2 else cfg.batch_size
dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = True
torch.backends.cudnn.allow_tf32 = True
tr = Trainer(cfg, dev)
img, lab = tr.synthetic_batch(batch)
for _ in range(3):
    tr.step(img, lab)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        tr.step(img, lab)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=90))
