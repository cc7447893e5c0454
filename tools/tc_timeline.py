"""Dump the clock64 phase timeline CTA 0 of the tensor-core FFT mixer records (debug hook
spectre_debug_tc_profile).  python tools/tc_timeline.py [--bwd] [B] [E]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vit_spectre_experiments_b200 import _lib  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
bwd = "--bwd" in sys.argv
B = int(args[0]) if args else 1024
E = int(args[1]) if len(args) > 1 else 384
H = W = 14
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH) if hasattr(_lib, "LIB_PATH") else lib
dev = torch.device("cuda:0")
T = H * W + 1
x = torch.randn(B, T, E, device=dev).bfloat16()
dy = torch.randn(B, T, E, device=dev).bfloat16()
f = torch.randn(H, W // 2 + 1, E, 2, device=dev)
y, dx, df = torch.empty_like(x), torch.empty_like(x), torch.empty_like(f)
st = torch.cuda.current_stream().cuda_stream
WHO, TILES, STAMPS = 6, 8, 16
prof = torch.zeros(2048, dtype=torch.int64, device=dev)   # [0, 768): phase stamps; [1024, ...): per-CTA wall clock
nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)


def run():
    if bwd:
        return lib.spectre_fftmix_bwd(dy.data_ptr(), x.data_ptr(), f.data_ptr(), dx.data_ptr(), df.data_ptr(),
                                      ws.data_ptr(), nws, B, H, W, E, 1, 1, st)
    return lib.spectre_fftmix_fwd(x.data_ptr(), f.data_ptr(), y.data_ptr(), B, H, W, E, 1, 1, st)


for _ in range(3):
    run()
torch.cuda.synchronize()
fn = lib.spectre_debug_tc_profile
fn.argtypes = [ctypes.c_void_p]
fn.restype = None
fn(prof.data_ptr())
run()
torch.cuda.synchronize()
fn(None)
pc = prof.cpu()
p = pc[:WHO * TILES * STAMPS].view(WHO, TILES, STAMPS)
t0 = int(p[0, 0, 14])
print(f"setup: start=0 setup_done={int(p[0,0,15])-t0} filter_loaded={int(p[0,0,13])-t0}")
names = {0: "epiWG0", 1: "epiWG1", 2: "mma", 3: "tma", 4: "aux4", 5: "aux5"}
for who in range(WHO):
    for t in range(TILES):
        row = p[who, t]
        vals = [int(v) - t0 if int(v) else None for v in row[:13]]
        if any(v is not None for v in vals):
            d = [vals[i] - vals[i - 1] if i and vals[i] is not None and vals[i - 1] is not None else 0 for i in range(len(vals))]
            print(f"{names[who]:7s} tile{t}: start {vals[0]} deltas " + " ".join(str(x) for x in d[1:]))
for t, name in ((0, "CTA 0"), (1, "last CTA")):
    q = p[4, t]
    ns, clk = int(q[1]) - int(q[0]), int(q[3]) - int(q[2])
    if ns > 0:
        print(f"{name}: wall {ns / 1e3:.1f} us, {clk} clk => SM clock {clk / ns * 1e3:.0f} MHz; started {int(q[0]) - int(p[4, 0, 0])} ns after CTA 0")
w = pc[1024:1024 + 2 * 148].view(148, 2)
live = w[:, 0] > 0
if live.any():
    t00 = int(w[live, 0].min())
    st = (w[live, 0] - t00).float() / 1e3
    en = (w[live, 1] - t00).float() / 1e3
    print(f"per-CTA wall clock ({int(live.sum())} CTAs): start min/max {st.min():.1f}/{st.max():.1f} us, "
          f"end min/median/max {en.min():.1f}/{en.median():.1f}/{en.max():.1f} us")
    worst = torch.argsort(en, descending=True)[:6]
    print("  latest CTAs:", [(int(i), round(float(en[i]), 1)) for i in worst])

