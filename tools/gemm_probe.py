"""Times the six GEMMs of one C4 encoder layer (rows = 2048*197, E = 384, hidden = 1536, bf16) in the operand layouts
torch can hand to cuBLAS, to see whether the layer's choice (addmm with a transposed weight view) is the fast one."""
import torch
dev = torch.device("cuda:0")
rows, E, hid = 2048 * 197, 384, 1536
bf = torch.bfloat16
x1 = torch.randn(rows, E, device=dev, dtype=bf); a1 = torch.randn(rows, hid, device=dev, dtype=bf)
dh1 = torch.randn(rows, hid, device=dev, dtype=bf); dh2 = torch.randn(rows, E, device=dev, dtype=bf)
W1 = torch.randn(hid, E, device=dev, dtype=bf); W3 = torch.randn(E, hid, device=dev, dtype=bf)
W1t, W3t = W1.t().contiguous(), W3.t().contiguous()
b1 = torch.randn(hid, device=dev, dtype=bf); b3 = torch.randn(E, device=dev, dtype=bf)
dx1 = torch.randn(rows, E, device=dev, dtype=bf); da1 = torch.randn(rows, hid, device=dev, dtype=bf)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def t(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts) // 2]
fl = 2 * rows * E * hid
cases = [
    ("h1 = addmm(b1, x1, W1.t())   [layer]", lambda: torch.addmm(b1, x1, W1.t())),
    ("h1 = addmm(b1, x1, W1t)", lambda: torch.addmm(b1, x1, W1t)),
    ("h1 = mm(x1, W1t) (no bias)", lambda: torch.mm(x1, W1t)),
    ("h2 = addmm(b3, a1, W3.t())   [layer]", lambda: torch.addmm(b3, a1, W3.t())),
    ("h2 = addmm(b3, a1, W3t)", lambda: torch.addmm(b3, a1, W3t)),
    ("d_a1 = mm(dh2, W3)           [layer]", lambda: torch.mm(dh2, W3)),
    ("d_a1 = mm(dh2, W3t.t())", lambda: torch.mm(dh2, W3t.t())),
    ("dw3 = mm(dh2.t(), a1)        [layer]", lambda: torch.mm(dh2.t(), a1)),
    ("dw3t = mm(a1.t(), dh2)", lambda: torch.mm(a1.t(), dh2)),
    ("d_x1.addmm_(dh1, W1)         [layer]", lambda: dx1.addmm_(dh1, W1)),
    ("d_x1.addmm_(dh1, W1t.t())", lambda: dx1.addmm_(dh1, W1t.t())),
    ("dw1 = mm(dh1.t(), x1)        [layer]", lambda: torch.mm(dh1.t(), x1)),
    ("dw1t = mm(x1.t(), dh1)", lambda: torch.mm(x1.t(), dh1)),
]
for name, fn in cases:
    ms = t(fn)
    print(f"{name:42s} {ms:.4f} ms  {fl / ms / 1e9:7.1f} TFLOP/s", flush=True)
