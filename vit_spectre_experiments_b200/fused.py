"""Fused row-wise tails of the encoder layer over the C ABI
(``spectre_norm_act_fwd`` / ``spectre_norm_act_bwd`` in ``include/spectre_b200.h``):

    out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + adaptive_avg_pool(skip) )

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
    """Returns ``(out_f32, out_bf16 or None)``; the bf16 twin is a second differentiable view of the same
    values (its incoming gradient is added to ``out``'s inside the backward kernel)."""

    @staticmethod
    def forward(ctx, in0, in1, skip, gamma, beta, eps, act, drop_p, seed, want_bf16):
        lib = _lib.load()
        if not in0.is_cuda:
            raise RuntimeError("norm_act: CUDA (sm_100a) only; there is no CPU fallback")
        if in0.dtype not in _DT:
            raise RuntimeError(f"norm_act: in0 must be float32 or bfloat16, got {in0.dtype}")
        N = in0.shape[-1]
        in0 = in0.contiguous()
        rows = in0.numel() // N
        in1 = None if in1 is None else in1.float().contiguous()
        skip = None if skip is None else skip.float().contiguous()
        gamma, beta = gamma.float().contiguous(), beta.float().contiguous()
        out = torch.empty(in0.shape, dtype=torch.float32, device=in0.device)
        out_bf16 = torch.empty(in0.shape, dtype=torch.bfloat16, device=in0.device) if want_bf16 else None
        mean = torch.empty(rows, dtype=torch.float32, device=in0.device)
        rstd = torch.empty_like(mean)
        d = _lib.NormActDesc()
        d.in0, d.in1, d.gamma, d.beta, d.skip = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta), _ptr(skip)
        d.out, d.out_bf16, d.mean, d.rstd = _ptr(out), _ptr(out_bf16), _ptr(mean), _ptr(rstd)
        d.rows, d.N, d.Ns = rows, N, (skip.shape[-1] if skip is not None else N)
        d.in0_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], int(act), float(eps), float(drop_p), int(seed)
        with torch.cuda.device(in0.device):
            with _timed("norm_act_fwd", (rows, N, in0.element_size())):
                rc = lib.spectre_norm_act_fwd(ctypes.byref(d), _stream_ptr(in0.device))
        _lib.check(rc, "spectre_norm_act_fwd")
        _lib.count_launches()
        ctx.save_for_backward(in0, in1, gamma, beta, mean, rstd)
        ctx.meta = (skip.shape if skip is not None else None, float(eps), int(act), float(drop_p), int(seed))
        ctx.need = (ctx.needs_input_grad[0], ctx.needs_input_grad[1] and in1 is not None)
        if out_bf16 is None:
            ctx.mark_non_differentiable()
            return out, None
        return out, out_bf16

    @staticmethod
    def backward(ctx, dout, dout_bf16):
        lib = _lib.load()
        in0, in1, gamma, beta, mean, rstd = ctx.saved_tensors
        skip_shape, eps, act, drop_p, seed = ctx.meta
        N = in0.shape[-1]
        rows = in0.numel() // N
        dev = in0.device
        if dout is None:
            dout = torch.zeros(in0.shape, dtype=torch.float32, device=dev)
        dout = dout.float().contiguous()
        if dout_bf16 is not None:
            dout_bf16 = dout_bf16.to(torch.bfloat16).contiguous()
        d_in0 = torch.empty_like(in0)
        # in1 is float32; when in0 is float32 too both get the same tensor
        d_in1 = None
        if in1 is not None and in0.dtype != torch.float32:
            d_in1 = torch.empty_like(in1)
        d_skip = torch.empty(skip_shape, dtype=torch.float32, device=dev) if skip_shape is not None else None
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(beta)
        d = _lib.NormActDesc()
        d.in0, d.in1, d.gamma, d.beta = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta)
        # `skip` itself is not read by the backward; a non-null pointer only selects the pooling mode
        d.skip = _ptr(d_skip)
        d.mean, d.rstd, d.dout, d.dout_bf16 = _ptr(mean), _ptr(rstd), _ptr(dout), _ptr(dout_bf16)
        d.d_in0, d.d_in1, d.d_skip, d.dgamma, d.dbeta = _ptr(d_in0), _ptr(d_in1), _ptr(d_skip), _ptr(dgamma), _ptr(dbeta)
        d.rows, d.N, d.Ns = rows, N, (skip_shape[-1] if skip_shape is not None else N)
        d.in0_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], act, eps, drop_p, seed
        with torch.cuda.device(dev):
            nws = lib.spectre_norm_act_bwd_workspace_bytes(rows, N)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
            d.workspace, d.workspace_bytes = ws.data_ptr(), nws
            with _timed("norm_act_bwd", (rows, N, in0.element_size())):
                rc = lib.spectre_norm_act_bwd(ctypes.byref(d), _stream_ptr(dev))
        _lib.check(rc, "spectre_norm_act_bwd")
        _lib.count_launches()
        g_in1 = None
        if in1 is not None:
            g_in1 = d_in0 if d_in1 is None else d_in1
        return d_in0, g_in1, d_skip, dgamma, dbeta, None, None, None, None, None


def norm_act(in0, gamma, beta, *, in1=None, skip=None, eps=1e-5, act="none", drop_p=0.0, seed=0, want_bf16=False):
    """``dropout(act(LayerNorm(in0 + in1) * gamma + beta) + adaptive_avg_pool(skip))`` in one CUDA pass.

    in0 ``[..., N]`` float32 / bfloat16; in1 ``[..., N]``, skip ``[..., Ns]`` float32 (or None).
    Returns ``(out_float32, out_bfloat16 | None)``.
    """
    return _NormAct.apply(in0, in1, skip, gamma, beta, eps, 1 if act == "gelu" else 0, drop_p, seed, want_bf16)
