"""Isolated timing of the 1-D Haar mixer kernels (dwt_token / dwt_embed) at the BASELINE.json token / channel counts:
CUDA events, L2 flushed between launches.  One JSON line per kernel; algorithmic bytes fwd 2*B*T*E*s + P*4,
bwd 3*B*T*E*s + 2*P*4 (P = weight elements)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vit_spectre_experiments_b200 import _lib  # noqa: E402

PEAK = 6552.3
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass

SHAPES = [("C3 8x8 E512 bf16 B1024", 1024, 64, 512, torch.bfloat16), ("C4 14x14 E384 bf16 B2048", 2048, 196, 384, torch.bfloat16),
          ("C2 7x7 E48 f32 B4096", 4096, 49, 48, torch.float32), ("8x8 E512 f32 B1024", 1024, 64, 512, torch.float32)]


def main():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = lambda: torch.cuda.current_stream(dev).cuda_stream
    for name, B, N, E, dt in SHAPES:
        for axis, aname in ((0, "token"), (1, "embed")):
            s, code = (4, 0) if dt == torch.float32 else (2, 1)
            x = torch.randn(B, N + 1, E, device=dev).to(dt)
            dy = torch.randn(B, N + 1, E, device=dev).to(dt)
            y, dx = torch.empty_like(x), torch.empty_like(x)
            w = torch.randn(*((2, (N + 1) // 2, E) if axis == 0 else (2, N, E // 2)), device=dev)
            dw = torch.empty_like(w)
            nws = lib.spectre_haar1d_bwd_workspace_bytes(B, N, E, axis)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
            fwd = lambda: _lib.check(lib.spectre_haar1d_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, N, E, 1, axis, code, st()), "fwd")
            bwd = lambda: _lib.check(lib.spectre_haar1d_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                                                            ws.data_ptr(), nws, B, N, E, 1, axis, code, st()), "bwd")
            P = w.numel() * 4
            for which, fn, nbytes in (("fwd", fwd, 2 * x.numel() * s + P), ("bwd", bwd, 3 * x.numel() * s + 2 * P)):
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(30):
                    flush.zero_()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(); fn(); b.record()
                    torch.cuda.synchronize()
                    ts.append(a.elapsed_time(b))
                ts.sort()
                med = ts[len(ts) // 2]
                print(json.dumps({"kernel": f"haar1d_{aname}_{which}", "shape": name, "ms_median": round(med, 4),
                                  "algo_MB": round(nbytes / 1e6, 2), "GBps_median": round(nbytes / med / 1e6, 1),
                                  "frac_of_measured_peak": round(nbytes / med / 1e6 / PEAK, 3)}), flush=True)


if __name__ == "__main__":
    main()
