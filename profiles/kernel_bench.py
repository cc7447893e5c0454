"""Isolated timing of every mixer kernel at the BASELINE.json shapes (CUDA events, inputs > L2
or L2 flushed between iterations).  Prints one JSON line per kernel:
    achieved GB/s = algorithmic bytes (fwd 2*T*E*s*B + F, bwd 3*T*E*s*B + 2F) / time.
Usage: python profiles/kernel_bench.py [--iters 20] [--only fft|dwt]
"""
import argparse
import json
import os
import re
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vit_spectre_experiments_b200 import _lib  # noqa: E402

PEAK = 6552.3
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass

ML_LEVELS = int(os.environ.get("SB200_KB_LEVELS", "3"))   # levels of the multi-level Haar shapes (names say J3)

SHAPES = [
    # name, kind, B, H, W, E, dtype
    ("C1 8x8 E512 f32 B8", "fft", 8, 8, 8, 512, torch.float32),
    ("C1-shape 8x8 E512 f32 B1024", "fft", 1024, 8, 8, 512, torch.float32),
    ("C2 7x7 E48 f32 B4096", "fft", 4096, 7, 7, 48, torch.float32),
    ("8x8 E512 bf16 B1024", "fft", 1024, 8, 8, 512, torch.bfloat16),
    ("C3 dwt 8x8 E512 bf16 B1024", "dwt", 1024, 8, 8, 512, torch.bfloat16),
    ("dwt 8x8 E512 f32 B1024", "dwt", 1024, 8, 8, 512, torch.float32),
    ("dwt 14x14 E384 bf16 B256", "dwt", 256, 14, 14, 384, torch.bfloat16),
    ("C3 dwtml J3 8x8 E512 bf16 B1024", "dwtml", 1024, 8, 8, 512, torch.bfloat16),
    ("dwtml J3 8x8 E512 f32 B1024", "dwtml", 1024, 8, 8, 512, torch.float32),
    ("C4 14x14 E384 bf16 B256", "fft", 256, 14, 14, 384, torch.bfloat16),
    ("C4-shape 14x14 E384 f32 B256", "fft", 256, 14, 14, 384, torch.float32),
    ("fp16-autocast ViT-S 14x14 E384 f32 B2048", "fft", 2048, 14, 14, 384, torch.float32),
    ("C5 14x14 E768 bf16 B128", "fft", 128, 14, 14, 768, torch.bfloat16),
    ("C4x4 14x14 E384 bf16 B1024", "fft", 1024, 14, 14, 384, torch.bfloat16),
    ("C4 bench 14x14 E384 bf16 B2048", "fft", 2048, 14, 14, 384, torch.bfloat16),
]


def time_ms(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()  # 256 MiB write: evicts the 126 MB L2
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="", help="substring match against fft | dwt | na (e.g. fftdwt)")
    ap.add_argument("--match", default="", help="regex searched in the shape name")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = lambda: torch.cuda.current_stream().cuda_stream
    for name, kind, B, H, W, E, dt in SHAPES:
        if (args.only and kind not in args.only) or (args.match and not re.search(args.match, name)):
            continue
        T = H * W + 1
        s = 4 if dt == torch.float32 else 2
        code = 0 if dt == torch.float32 else 1
        x = torch.randn(B, T, E, device=dev).to(dt)
        dy = torch.randn(B, T, E, device=dev).to(dt)
        y = torch.empty_like(x)
        dx = torch.empty_like(x)
        if kind == "fft":
            p = torch.randn(H, W // 2 + 1, E, 2, device=dev)
            nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
        elif kind == "dwtml":
            p = torch.randn(H, W, E, device=dev)
            nws = lib.spectre_dwtmix_ml_bwd_workspace_bytes(B, H, W, E)
        else:
            p = torch.randn(4, H // 2, W // 2, E, device=dev)
            nws = lib.spectre_dwtmix_bwd_workspace_bytes(B, H, W, E)
        dp = torch.empty_like(p)
        ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
        F = p.numel() * 4
        if kind == "fft":
            fwd = lambda: _lib.check(lib.spectre_fftmix_fwd(x.data_ptr(), p.data_ptr(), y.data_ptr(), B, H, W, E, 1, code, st()), "fwd")
            bwd = lambda: _lib.check(lib.spectre_fftmix_bwd(dy.data_ptr(), x.data_ptr(), p.data_ptr(), dx.data_ptr(), dp.data_ptr(), ws.data_ptr(), nws, B, H, W, E, 1, code, st()), "bwd")
        elif kind == "dwtml":
            fwd = lambda: _lib.check(lib.spectre_dwtmix_ml_fwd(x.data_ptr(), p.data_ptr(), y.data_ptr(), B, H, W, E, 1, ML_LEVELS, code, st()), "fwd")
            bwd = lambda: _lib.check(lib.spectre_dwtmix_ml_bwd(dy.data_ptr(), x.data_ptr(), p.data_ptr(), dx.data_ptr(), dp.data_ptr(), ws.data_ptr(), nws, B, H, W, E, 1, ML_LEVELS, code, st()), "bwd")
        else:
            fwd = lambda: _lib.check(lib.spectre_dwtmix_fwd(x.data_ptr(), p.data_ptr(), y.data_ptr(), B, H, W, E, 1, code, st()), "fwd")
            bwd = lambda: _lib.check(lib.spectre_dwtmix_bwd(dy.data_ptr(), x.data_ptr(), p.data_ptr(), dx.data_ptr(), dp.data_ptr(), ws.data_ptr(), nws, B, H, W, E, 1, code, st()), "bwd")
        for which, fn, nbytes in (("fwd", fwd, 2 * B * T * E * s + F), ("bwd", bwd, 3 * B * T * E * s + 2 * F)):
            med, best = time_ms(fn, args.iters, flush)
            print(json.dumps({
                "kernel": f"{kind}mix_{which}", "shape": name, "ms_median": round(med, 4), "ms_best": round(best, 4),
                "algo_MB": round(nbytes / 1e6, 2), "GBps_median": round(nbytes / med / 1e6, 1),
                "GBps_best": round(nbytes / best / 1e6, 1), "frac_of_measured_peak": round(nbytes / med / 1e6 / PEAK, 3),
                "workspace_MB": round(nws / 1e6, 2),
            }), flush=True)

    if args.only == "" or "na" in args.only:
        norm_act_bench(args, dev, flush)


# the row-wise tails of one encoder layer as fused.encoder_layer calls them (bf16 stream unless noted)
#   name, rows, N, Ns (skip width, 0 = none), in1, act, drop_p, dtype, pool_k (fwd out_pool / bwd d_out_pool), dout2, colsum
def _layer_shapes(tag, rows, E, hid, dt=torch.bfloat16):
    return [
        (f"{tag} LN1(mix)+x N{E}", rows, E, E, False, 0, 0.0, dt, 0, True, False),
        (f"{tag} GELU(LN(h1))+up+pool N{hid}", rows, hid, E, False, 1, 0.001, dt, hid // E, False, True),
        (f"{tag} GELU(LN(h2))+pooled N{E}", rows, E, E, False, 1, 0.001, dt, 0, False, True),
        (f"{tag} LN2(x1+ff) N{E}", rows, E, 0, True, 0, 0.0, dt, 0, True, False),
    ]


NA_SHAPES = (
    _layer_shapes("C4 B256", 256 * 197, 384, 1536)
    + _layer_shapes("C4 B2048", 2048 * 197, 384, 1536)
    + _layer_shapes("C5 B512", 512 * 197, 768, 3072)
    + [
        ("C4 B2048 GELU(LN(h2))+down N384 (r1 layout)", 2048 * 197, 384, 1536, False, 1, 0.001, torch.bfloat16, 0, False, True),
        ("C4 B256 GELU(LN(h1))+up N1536 f32", 256 * 197, 1536, 384, False, 1, 0.001, torch.float32, 0, False, True),
        ("C3 GELU(LN(h1))+general 512->768", 1024 * 65, 768, 512, False, 1, 0.001, torch.bfloat16, 0, False, True),
        ("C3 GELU(LN(h2))+general 768->512", 1024 * 65, 512, 768, False, 1, 0.001, torch.bfloat16, 0, False, True),
        ("C3 LN1(mix)+x N512", 1024 * 65, 512, 512, False, 0, 0.0, torch.bfloat16, 0, True, False),
        ("C2 GELU(LN(h1))+general 48->256 f32", 4096 * 50, 256, 48, False, 1, 0.001, torch.float32, 0, False, True),
        ("C2 GELU(LN(h2))+general 256->48 f32", 4096 * 50, 48, 256, False, 1, 0.001, torch.float32, 0, False, True),
    ]
)


def norm_act_bench(args, dev, flush):
    from vit_spectre_experiments_b200.functional import TIMER
    from vit_spectre_experiments_b200.fused import na_bwd, na_fwd

    for name, rows, N, Ns, has_in1, act, p, dt, pool_k, has_d2, colsum in NA_SHAPES:
        if args.match and not re.search(args.match, name):
            continue
        rnd = lambda *s: torch.randn(*s, device=dev).to(dt)   # noqa: E731
        in0 = rnd(rows, N)
        in1 = rnd(rows, N) if has_in1 else None
        skip = rnd(rows, Ns) if Ns else None
        gamma, beta = torch.ones(N, device=dev), torch.zeros(N, device=dev)
        dout = rnd(rows, N)
        dout2 = rnd(rows, N) if has_d2 else None
        dpool = rnd(rows, N // pool_k) if pool_k else None
        res = na_fwd(in0, gamma, beta, in1, skip, act=act, drop_p=p, seed=11, pool_k=pool_k)
        mean, rstd = res[1], res[2]

        def fwd():
            na_fwd(in0, gamma, beta, in1, skip, act=act, drop_p=p, seed=11, pool_k=pool_k)

        def bwd():
            na_bwd(dout, in0, gamma, beta, mean, rstd, in1, (rows, Ns) if Ns else None, act=act, drop_p=p, seed=11,
                   dout2=dout2, colsum=colsum, d_out_pool=dpool, pool_k=pool_k)

        for which, fn in (("fwd", fwd), ("bwd", bwd)):
            ts = []
            nbytes = 0
            for it in range(args.iters + 3):
                flush.zero_()
                TIMER.reset(); TIMER.enabled = True
                fn()
                TIMER.enabled = False
                (n, ms, meta, main_ms), = TIMER.summary().values()
                nbytes = meta[3]
                if it >= 3:
                    ts.append(ms)
            ts.sort()
            med, best = ts[len(ts) // 2], ts[0]
            print(json.dumps({
                "kernel": f"norm_act_{which}", "shape": name, "ms_median": round(med, 4), "ms_best": round(best, 4),
                "algo_MB": round(nbytes / 1e6, 2), "GBps_median": round(nbytes / med / 1e6, 1),
                "frac_of_measured_peak": round(nbytes / med / 1e6 / PEAK, 3),
            }), flush=True)
        del in0, in1, skip, dout, dout2, dpool, res
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
