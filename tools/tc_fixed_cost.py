"""Fixed vs per-tile cost of the tensor-core mixers: time fwd / bwd at batch sizes that give every CTA 0, 1, 2, 4, ...
tiles (14x14, E=384, bf16).  python tools/tc_fixed_cost.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vit_spectre_experiments_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda:0")
H = W = 14
E = 384
T = H * W + 1
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
f = torch.randn(H, W // 2 + 1, E, 2, device=dev)
df = torch.empty_like(f)
for B in (2, 49, 98, 196, 392, 784, 1024, 2048):
    x = torch.randn(B, T, E, device=dev).bfloat16()
    dy = torch.randn(B, T, E, device=dev).bfloat16()
    y, dx = torch.empty_like(x), torch.empty_like(x)
    nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
    ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
    fwd = lambda: lib.spectre_fftmix_fwd(x.data_ptr(), f.data_ptr(), y.data_ptr(), B, H, W, E, 1, 1, st)
    bwd = lambda: lib.spectre_fftmix_bwd(dy.data_ptr(), x.data_ptr(), f.data_ptr(), dx.data_ptr(), df.data_ptr(),
                                         ws.data_ptr(), nws, B, H, W, E, 1, 1, st)
    out = []
    for fn in (fwd, bwd):
        ts = []
        for _ in range(12):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        ts.sort()
        out.append(ts[len(ts) // 2])
    print(f"B={B:5d}  fwd {out[0]:7.1f} us ({(B * 3 + 146) // 147} tiles/CTA)   bwd {out[1]:7.1f} us ({((B + 1) // 2 * 6 + 143) // 144} tiles/CTA)")
