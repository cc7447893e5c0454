"""Fused row-wise tails of the encoder layer over the C ABI
(``spectre_norm_act_fwd`` / ``spectre_norm_act_bwd`` in ``include/spectre_b200.h``):

    out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + adaptive_avg_pool(skip) )

Every activation operand of a call has one dtype: float32, or bfloat16 for the bf16 training stream (the
encoder keeps its residual stream in bf16 under bf16 autocast -- half the HBM traffic of autocast's
fp32 LayerNorm outputs; statistics and arithmetic stay fp32).

One CUDA pass each way replaces the LayerNorm / GELU / AdaptiveAvgPool1d / add / dropout launches of
``SpectreEncoderLayer.forward`` (``spectre_vit/models/spectre/spectre.py:65-73``) and
``SpectreLinear.forward`` (``spectre_vit/models/spectre/layers.py:95-101``).  CUDA only, no fallback:
the torch composition of the same modules stays the (CPU-capable) reference path in ``model.py``.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .functional import TIMER, _stream_ptr, _timed

_DT = {torch.float32: 0, torch.bfloat16: 1}


def _ptr(t):
    return None if t is None else t.data_ptr()


class _NormAct(torch.autograd.Function):
    """All activation operands share ``in0``'s dtype (float32, or bfloat16 for the bf16 stream)."""

    @staticmethod
    def forward(ctx, in0, in1, skip, gamma, beta, eps, act, drop_p, seed):
        lib = _lib.load()
        if not in0.is_cuda:
            raise RuntimeError("norm_act: CUDA (sm_100a) only; there is no CPU fallback")
        if in0.dtype not in _DT:
            raise RuntimeError(f"norm_act: in0 must be float32 or bfloat16, got {in0.dtype}")
        N = in0.shape[-1]
        in0 = in0.contiguous()
        rows = in0.numel() // N
        in1 = None if in1 is None else in1.to(in0.dtype).contiguous()
        skip = None if skip is None else skip.to(in0.dtype).contiguous()
        gamma, beta = gamma.float().contiguous(), beta.float().contiguous()
        out = torch.empty_like(in0)
        mean = torch.empty(rows, dtype=torch.float32, device=in0.device)
        rstd = torch.empty_like(mean)
        d = _lib.NormActDesc()
        d.in0, d.in1, d.gamma, d.beta, d.skip = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta), _ptr(skip)
        d.out, d.mean, d.rstd = _ptr(out), _ptr(mean), _ptr(rstd)
        d.rows, d.N, d.Ns = rows, N, (skip.shape[-1] if skip is not None else N)
        d.act_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], int(act), float(eps), float(drop_p), int(seed)
        with torch.cuda.device(in0.device):
            with _timed("norm_act_fwd", (rows, N, in0.element_size())):
                rc = lib.spectre_norm_act_fwd(ctypes.byref(d), _stream_ptr(in0.device))
        _lib.check(rc, "spectre_norm_act_fwd")
        _lib.count_launches()
        ctx.save_for_backward(in0, in1, gamma, beta, mean, rstd)
        ctx.meta = (skip.shape if skip is not None else None, float(eps), int(act), float(drop_p), int(seed))
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.load()
        in0, in1, gamma, beta, mean, rstd = ctx.saved_tensors
        skip_shape, eps, act, drop_p, seed = ctx.meta
        N = in0.shape[-1]
        rows = in0.numel() // N
        dev = in0.device
        dout = dout.to(in0.dtype).contiguous()
        d_in0 = torch.empty_like(in0)
        d_skip = torch.empty(skip_shape, dtype=in0.dtype, device=dev) if skip_shape is not None else None
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(beta)
        d = _lib.NormActDesc()
        d.in0, d.in1, d.gamma, d.beta = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta)
        d.mean, d.rstd, d.dout = _ptr(mean), _ptr(rstd), _ptr(dout)
        d.d_in0, d.d_skip, d.dgamma, d.dbeta = _ptr(d_in0), _ptr(d_skip), _ptr(dgamma), _ptr(dbeta)
        d.rows, d.N, d.Ns = rows, N, (skip_shape[-1] if skip_shape is not None else N)
        d.act_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], act, eps, drop_p, seed
        with torch.cuda.device(dev):
            nws = lib.spectre_norm_act_bwd_workspace_bytes(rows, N)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
            d.workspace, d.workspace_bytes = ws.data_ptr(), nws
            with _timed("norm_act_bwd", (rows, N, in0.element_size())):
                rc = lib.spectre_norm_act_bwd(ctypes.byref(d), _stream_ptr(dev))
        _lib.check(rc, "spectre_norm_act_bwd")
        _lib.count_launches()
        # d(in0 + in1): the same values for both addends
        return d_in0, (d_in0 if in1 is not None else None), d_skip, dgamma, dbeta, None, None, None, None


def norm_act(in0, gamma, beta, *, in1=None, skip=None, eps=1e-5, act="none", drop_p=0.0, seed=0):
    """``dropout(act(LayerNorm(in0 + in1) * gamma + beta) + adaptive_avg_pool(skip))`` in one CUDA pass.

    in0 ``[..., N]`` float32 / bfloat16; in1 ``[..., N]``, skip ``[..., Ns]`` (cast to in0's dtype if they
    differ).  Returns ``out`` in in0's dtype; statistics and arithmetic are float32.
    """
    return _NormAct.apply(in0, in1, skip, gamma, beta, eps, 1 if act == "gelu" else 0, drop_p, seed)
