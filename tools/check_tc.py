"""Quick GPU check of the bf16 FFT-mixer kernels against the oracle + isolated timing.

    python tools/check_tc.py [--bwd] [--time]

Development aid (the judged parity tests are tests/test_gpu_parity.py; the judged timings are
bench.py and profiles/kernel_bench.py).  Uses oracle/ only as the checker.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import mixers_ref as O  # noqa: E402
from vit_spectre_experiments_b200 import _lib  # noqa: E402

CASES = [
    (4, 14, 14, 384, 1), (2, 14, 14, 768, 1), (5, 8, 8, 512, 1), (3, 7, 7, 64, 1), (300, 14, 14, 64, 1),
    (3, 14, 14, 192, 0), (2, 8, 8, 64, 0), (7, 7, 7, 128, 1), (2, 14, 14, 200, 1), (1, 8, 8, 8, 1),
]


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bwd", action="store_true")
    ap.add_argument("--time", action="store_true")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    st = lambda: torch.cuda.current_stream().cuda_stream
    worst = 0.0
    for B, H, W, E, cls in CASES:
        g = torch.Generator().manual_seed(B * 1000 + E)
        T = H * W + cls
        x = torch.randn(B, T, E, generator=g).bfloat16()
        dy = torch.randn(B, T, E, generator=g).bfloat16()
        filt = torch.randn(H, W // 2 + 1, E, 2, generator=g)
        xd, dyd, fd = x.to(dev), dy.to(dev), filt.to(dev)
        y = torch.full_like(xd, float("nan"))
        rc = lib.spectre_fftmix_fwd(xd.data_ptr(), fd.data_ptr(), y.data_ptr(), B, H, W, E, cls, 1, st())
        _lib.check(rc, "fwd")
        torch.cuda.synchronize()
        xr, dyr = x.float(), dy.float()
        if not cls:
            xr = torch.cat([torch.zeros(B, 1, E), xr], 1)
            dyr = torch.cat([torch.zeros(B, 1, E), dyr], 1)
        y_ref = O.fft_mix_ref(xr, filt, H, W)
        if not cls:
            y_ref = y_ref[:, 1:]
        e_y = rel(y, y_ref.bfloat16())
        msg = f"B={B:4d} {H}x{W} E={E:4d} cls={cls}  fwd rel={e_y:.2e}"
        worst = max(worst, e_y if e_y == e_y else 1e9)
        if args.bwd:
            dx = torch.full_like(xd, float("nan"))
            df = torch.full_like(fd, float("nan"))
            nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
            rc = lib.spectre_fftmix_bwd(dyd.data_ptr(), xd.data_ptr(), fd.data_ptr(), dx.data_ptr(), df.data_ptr(),
                                        ws.data_ptr(), nws, B, H, W, E, cls, 1, st())
            _lib.check(rc, "bwd")
            torch.cuda.synchronize()
            dx_ref, df_ref = O.fft_mix_bwd_ref(dyr, xr, filt, H, W)
            if not cls:
                dx_ref = dx_ref[:, 1:]
            e_dx, e_df = rel(dx, dx_ref.bfloat16()), rel(df, df_ref)
            msg += f"  dx rel={e_dx:.2e}  dfilt rel={e_df:.2e}"
            for e in (e_dx, e_df):
                worst = max(worst, e if e == e else 1e9)
        print(msg, flush=True)
    print(f"worst rel err {worst:.3e}  ({'OK' if worst <= 3e-2 else 'FAIL'})", flush=True)

    if args.time:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for B, H, W, E in [(256, 14, 14, 384), (1024, 14, 14, 384), (128, 14, 14, 768), (1024, 8, 8, 512)]:
            T = H * W + 1
            x = torch.randn(B, T, E, device=dev).bfloat16()
            dy = torch.randn(B, T, E, device=dev).bfloat16()
            f = torch.randn(H, W // 2 + 1, E, 2, device=dev)
            y, dx, df = torch.empty_like(x), torch.empty_like(x), torch.empty_like(f)
            nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
            fwd = lambda: lib.spectre_fftmix_fwd(x.data_ptr(), f.data_ptr(), y.data_ptr(), B, H, W, E, 1, 1, st())
            bwd = lambda: lib.spectre_fftmix_bwd(dy.data_ptr(), x.data_ptr(), f.data_ptr(), dx.data_ptr(), df.data_ptr(),
                                                 ws.data_ptr(), nws, B, H, W, E, 1, 1, st())
            fns = [("fwd", fwd, 2 * B * T * E * 2 + f.numel() * 4)]
            if args.bwd:
                fns.append(("bwd", bwd, 3 * B * T * E * 2 + 2 * f.numel() * 4))
            for name, fn, nbytes in fns:
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(15):
                    flush.zero_()
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record(); fn(); e.record()
                    torch.cuda.synchronize()
                    ts.append(s.elapsed_time(e))
                ts.sort()
                med = ts[len(ts) // 2]
                print(f"time {name} B={B} {H}x{W} E={E}: median {med*1e3:.1f} us best {ts[0]*1e3:.1f} us  "
                      f"{nbytes/med/1e6:.0f} GB/s ({nbytes/med/1e6/6553:.3f} of measured peak)", flush=True)
    return 0 if worst <= 3e-2 else 1


if __name__ == "__main__":
    sys.exit(main())
