"""Fused row-wise tails of the encoder layer over the C ABI
(``spectre_norm_act_fwd`` / ``spectre_norm_act_bwd`` in ``include/spectre_b200.h``):

    out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + adaptive_avg_pool(skip) )

Every activation operand of a call has one dtype: float32, or bfloat16 for the bf16 training stream (the
encoder keeps its residual stream in bf16 under bf16 autocast -- half the HBM traffic of autocast's
fp32 LayerNorm outputs; statistics and arithmetic stay fp32).

One CUDA pass each way replaces the LayerNorm / GELU / AdaptiveAvgPool1d / add / dropout launches of
``SpectreEncoderLayer.forward`` (``spectre_vit/models/spectre/spectre.py:65-73``) and
``SpectreLinear.forward`` (``spectre_vit/models/spectre/layers.py:95-101``).  Two levels:

* :func:`norm_act` -- one tail as an autograd function;
* :func:`encoder_layer` -- the whole layer (mixer, four tails, two GEMMs) as ONE autograd node with a
  hand-scheduled backward: gradient sums are folded into kernels that run anyway (second ``dout``
  operand of the tail backward, ``addmm`` epilogue of the input-gradient GEMMs) and the Linear bias
  gradients come out of the tail backward as column sums, so no stand-alone add / reduce launches remain.

CUDA only, no fallback: the torch composition of the same modules stays the (CPU-capable) reference path
in ``model.py``.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .functional import _stream_ptr, _timed, dwtmix_bwd_raw, dwtmix_fwd_raw, fftmix_bwd_raw, fftmix_fwd_raw

_DT = {torch.float32: 0, torch.bfloat16: 1}


def _ptr(t):
    return None if t is None else t.data_ptr()


# ------------------------------------------------------------------ raw (non-autograd) kernel calls
def na_fwd(in0, gamma, beta, in1=None, skip=None, eps=1e-5, act=0, drop_p=0.0, seed=0, reverse=False, pool_k=0):
    """One forward tail.  Returns ``(out, mean, rstd)``; operands must already share in0's dtype.
    ``pool_k`` in (2, 4, 8): also returns ``out_pool [rows, N / pool_k]`` (mean of every ``pool_k`` consecutive columns
    of ``out``) as a fourth value."""
    lib = _lib.load()
    if not in0.is_cuda:
        raise RuntimeError("norm_act: CUDA (sm_100a) only; there is no CPU fallback")
    if in0.dtype not in _DT:
        raise RuntimeError(f"norm_act: in0 must be float32 or bfloat16, got {in0.dtype}")
    N = in0.shape[-1]
    rows = in0.numel() // N
    out = torch.empty_like(in0)
    mean = torch.empty(rows, dtype=torch.float32, device=in0.device)
    rstd = torch.empty_like(mean)
    d = _lib.NormActDesc()
    d.in0, d.in1, d.gamma, d.beta, d.skip = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta), _ptr(skip)
    d.out, d.mean, d.rstd = _ptr(out), _ptr(mean), _ptr(rstd)
    d.rows, d.N, d.Ns = rows, N, (skip.shape[-1] if skip is not None else N)
    d.act_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], int(act), float(eps), float(drop_p), int(seed)
    d.reverse = int(bool(reverse))
    d.seed_dev = _lib.seed_dev_ptr() if drop_p > 0 else None
    out_pool = None
    if pool_k:
        out_pool = torch.empty(rows, N // pool_k, dtype=in0.dtype, device=in0.device)
        d.out_pool, d.pool_k = _ptr(out_pool), int(pool_k)
    nbytes = (in0.numel() * (2 + (in1 is not None)) + (skip.numel() if skip is not None else 0)
              + (out_pool.numel() if pool_k else 0)) * in0.element_size()
    with torch.cuda.device(in0.device):
        tag = f"norm_act_fwd:N{N}:Ns{d.Ns if skip is not None else 0}:act{int(act)}:in1{int(in1 is not None)}"
        with _timed(tag, (rows, N, in0.element_size(), nbytes)):
            rc = lib.spectre_norm_act_fwd(ctypes.byref(d), _stream_ptr(in0.device))
    _lib.check(rc, "spectre_norm_act_fwd")
    _lib.count_launches()
    if pool_k:
        return out, mean, rstd, out_pool
    return out, mean, rstd


def na_bwd(dout, in0, gamma, beta, mean, rstd, in1=None, skip_shape=None, eps=1e-5, act=0, drop_p=0.0, seed=0,
           dout2=None, colsum=False, reverse=False, d_out_pool=None, pool_k=0):
    """One backward tail.  Returns ``(d_in0, d_skip | None, dgamma, dbeta, dcolsum | None)``; ``d_in0`` is also
    the gradient of ``in1``; ``dcolsum`` = column sums of ``d_in0`` (fp32).  ``d_out_pool [rows, N / pool_k]``: gradient
    that arrived through the forward's pooled copy (added to ``dout`` inside the kernel)."""
    lib = _lib.load()
    N = in0.shape[-1]
    rows = in0.numel() // N
    dev = in0.device
    d_in0 = torch.empty_like(in0)
    d_skip = torch.empty(skip_shape, dtype=in0.dtype, device=dev) if skip_shape is not None else None
    dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(beta)
    dcol = torch.empty_like(gamma) if colsum else None
    d = _lib.NormActDesc()
    d.in0, d.in1, d.gamma, d.beta = _ptr(in0), _ptr(in1), _ptr(gamma), _ptr(beta)
    d.mean, d.rstd, d.dout, d.dout2 = _ptr(mean), _ptr(rstd), _ptr(dout), _ptr(dout2)
    d.d_in0, d.d_skip, d.dgamma, d.dbeta, d.dcolsum = _ptr(d_in0), _ptr(d_skip), _ptr(dgamma), _ptr(dbeta), _ptr(dcol)
    d.rows, d.N, d.Ns = rows, N, (skip_shape[-1] if skip_shape is not None else N)
    d.act_dtype, d.act, d.eps, d.drop_p, d.seed = _DT[in0.dtype], int(act), float(eps), float(drop_p), int(seed)
    d.reverse = int(bool(reverse))
    d.seed_dev = _lib.seed_dev_ptr() if drop_p > 0 else None
    if d_out_pool is not None:
        d.d_out_pool, d.pool_k = _ptr(d_out_pool), int(pool_k)
    nbytes = (in0.numel() * (3 + (in1 is not None) + (dout2 is not None)) + (d_skip.numel() if d_skip is not None else 0)
              + (d_out_pool.numel() if d_out_pool is not None else 0)) * in0.element_size()
    with torch.cuda.device(dev):
        nws = lib.spectre_norm_act_bwd_workspace_bytes(rows, N)
        ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
        d.workspace, d.workspace_bytes = ws.data_ptr(), nws
        tag = f"norm_act_bwd:N{N}:Ns{d.Ns if d_skip is not None else 0}:act{int(act)}:in1{int(in1 is not None)}"
        with _timed(tag, (rows, N, in0.element_size(), nbytes)):
            rc = lib.spectre_norm_act_bwd(ctypes.byref(d), _stream_ptr(dev))
    _lib.check(rc, "spectre_norm_act_bwd")
    _lib.count_launches()
    return d_in0, d_skip, dgamma, dbeta, dcol


# ------------------------------------------------------------------ one tail as an autograd function
class _NormAct(torch.autograd.Function):
    """All activation operands share ``in0``'s dtype (float32, or bfloat16 for the bf16 stream)."""

    @staticmethod
    def forward(ctx, in0, in1, skip, gamma, beta, eps, act, drop_p, seed):
        in0 = in0.contiguous()
        in1 = None if in1 is None else in1.to(in0.dtype).contiguous()
        skip = None if skip is None else skip.to(in0.dtype).contiguous()
        gamma, beta = gamma.float().contiguous(), beta.float().contiguous()
        out, mean, rstd = na_fwd(in0, gamma, beta, in1, skip, eps, act, drop_p, seed)
        ctx.save_for_backward(in0, in1, gamma, beta, mean, rstd)
        ctx.meta = (skip.shape if skip is not None else None, float(eps), int(act), float(drop_p), int(seed))
        return out

    @staticmethod
    def backward(ctx, dout):
        in0, in1, gamma, beta, mean, rstd = ctx.saved_tensors
        skip_shape, eps, act, drop_p, seed = ctx.meta
        dout = dout.to(in0.dtype).contiguous()
        d_in0, d_skip, dgamma, dbeta, _ = na_bwd(dout, in0, gamma, beta, mean, rstd, in1, skip_shape, eps, act,
                                                 drop_p, seed)
        # d(in0 + in1): the same values for both addends
        return d_in0, (d_in0 if in1 is not None else None), d_skip, dgamma, dbeta, None, None, None, None


def norm_act(in0, gamma, beta, *, in1=None, skip=None, eps=1e-5, act="none", drop_p=0.0, seed=0):
    """``dropout(act(LayerNorm(in0 + in1) * gamma + beta) + adaptive_avg_pool(skip))`` in one CUDA pass.

    in0 ``[..., N]`` float32 / bfloat16; in1 ``[..., N]``, skip ``[..., Ns]`` (cast to in0's dtype if they
    differ).  Returns ``out`` in in0's dtype; statistics and arithmetic are float32.
    """
    return _NormAct.apply(in0, in1, skip, gamma, beta, eps, 1 if act == "gelu" else 0, drop_p, seed)


# ------------------------------------------------------------------ the whole encoder layer as one node
class _EncoderLayer(torch.autograd.Function):
    """``x1 = LN1(mix(x)) + x ; x2 = LN2(x1 + SL3(drop(SL1(x1))))`` with ``SL(u) = GELU(LN(Linear(u))) + pool(u)``
    (spectre.py:65-73, layers.py:95-101).  Seven kernels + two GEMMs forward; backward is scheduled by hand."""

    # Gradient hand-over between stacked layers.  The input gradient of a layer is the sum of two tensors (the identity
    # skip of LN1 and the mixer backward); the layer below can add them for free inside its first backward kernel
    # (``dout2``).  So the node has TWO outputs -- the result and a second view of the same buffer -- and takes the pair
    # of the layer below as ``x`` / ``x_twin``: autograd then routes the two addends separately (grad of x = skip
    # part, grad of x_twin = mixer part) and no stand-alone add runs between layers (it was 1.7 % of the C4 step).
    # Only ``x`` is read in forward; with ``x_twin=None`` (first layer, stand-alone use) the backward adds itself.
    @staticmethod
    def forward(ctx, x, x_twin, mix_param, n1w, n1b, w1, b1, l1w, l1b, w3, b3, l3w, l3b, n2w, n2b, cfg):
        kind, H, W, has_cls, eps, p1, p2, seed1, seed2 = cfg
        ctx.set_materialize_grads(False)
        ctx.has_twin = x_twin is not None
        x = x.contiguous()
        dt = x.dtype
        B, T, E = x.shape
        rows = B * T
        f32 = lambda t: t.detach().float().contiguous()   # noqa: E731
        n1w, n1b, l1w, l1b, l3w, l3b, n2w, n2b = map(f32, (n1w, n1b, l1w, l1b, l3w, l3b, n2w, n2b))
        mp = f32(mix_param)
        w1c, w3c = w1.detach().to(dt), w3.detach().to(dt)          # GEMM operands in the stream dtype
        b1c, b3c = b1.detach().to(dt), b3.detach().to(dt)
        m = (fftmix_fwd_raw if kind == "fft" else dwtmix_fwd_raw)(x, mp, H, W, has_cls)
        x2d, m2d = x.view(rows, E), m.view(rows, E)
        x1, mean1, rstd1 = na_fwd(m2d, n1w, n1b, skip=x2d, eps=eps[0])
        h1 = torch.addmm(b1c, x1, w1c.t())
        # hidden = k * E with k in (2, 4, 8): linear3's skip AdaptiveAvgPool1d(E)(a1) is the mean of every k consecutive
        # columns -- the kernel that writes a1 emits it (out_pool), so the second tail reads E columns per row, not k * E
        hid = w1.shape[0]
        pk = hid // E if (hid % E == 0 and hid // E in (2, 4, 8) and hid % (8 * (hid // E)) == 0) else 0
        ctx.pool_k = pk
        if pk:
            a1, mean2, rstd2, a1p = na_fwd(h1, l1w, l1b, skip=x1, eps=eps[1], act=1, drop_p=p1, seed=seed1, pool_k=pk)
        else:
            a1, mean2, rstd2 = na_fwd(h1, l1w, l1b, skip=x1, eps=eps[1], act=1, drop_p=p1, seed=seed1)
            a1p = a1
        h2 = torch.addmm(b3c, a1, w3c.t())
        ff, mean3, rstd3 = na_fwd(h2, l3w, l3b, skip=a1p, eps=eps[2], act=1, drop_p=p2, seed=seed2)
        del a1p
        x2, mean4, rstd4 = na_fwd(x1, n2w, n2b, in1=ff, eps=eps[3])
        ctx.save_for_backward(x, mp, m, x1, h1, a1, h2, ff, w1c, w3c, n1w, n1b, l1w, l1b, l3w, l3b, n2w, n2b,
                              mean1, rstd1, mean2, rstd2, mean3, rstd3, mean4, rstd4)
        ctx.cfg = cfg
        ctx.param_dtypes = (w1.dtype, b1.dtype, w3.dtype, b3.dtype)
        return x2.view(B, T, E), x2.view(B, T, E)

    @staticmethod
    def backward(ctx, d_x2, d_x2b):
        (x, mp, m, x1, h1, a1, h2, ff, w1c, w3c, n1w, n1b, l1w, l1b, l3w, l3b, n2w, n2b,
         mean1, rstd1, mean2, rstd2, mean3, rstd3, mean4, rstd4) = ctx.saved_tensors
        kind, H, W, has_cls, eps, p1, p2, seed1, seed2 = ctx.cfg
        B, T, E = x.shape
        rows = B * T
        hid = h1.shape[-1]
        if d_x2 is None:
            d_x2, d_x2b = d_x2b, None
        d_x2 = d_x2.to(x.dtype).contiguous().view(rows, E)
        if d_x2b is not None:
            d_x2b = d_x2b.to(x.dtype).contiguous().view(rows, E)
        # x2 = LN2(x1 + ff): one gradient for both addends; the two parts of the upstream gradient are summed in the kernel
        d_z, _, dn2w, dn2b, _ = na_bwd(d_x2, x1, n2w, n2b, mean4, rstd4, in1=ff, eps=eps[3], dout2=d_x2b)
        # ff = GELU(LN(h2)) + pool_down(a1)
        pk = ctx.pool_k
        if pk:
            # the skip gradient stays [rows, E] (gradient of the pooled copy); the first tail's backward spreads it
            d_h2, d_a1p, dl3w, dl3b, db3 = na_bwd(d_z, h2, l3w, l3b, mean3, rstd3, skip_shape=(rows, E), eps=eps[2], act=1,
                                                  drop_p=p2, seed=seed2, colsum=True)
            dw3 = torch.mm(d_h2.t(), a1)
            d_a1 = torch.mm(d_h2, w3c)
        else:
            d_h2, d_a1, dl3w, dl3b, db3 = na_bwd(d_z, h2, l3w, l3b, mean3, rstd3, skip_shape=(rows, hid), eps=eps[2],
                                                 act=1, drop_p=p2, seed=seed2, colsum=True)
            dw3 = torch.mm(d_h2.t(), a1)
            d_a1.addmm_(d_h2, w3c)                                 # skip path + GEMM path in the GEMM epilogue (in place)
            d_a1p = None
        # a1 = drop(GELU(LN(h1)) + pool_up(x1))
        d_h1, d_x1, dl1w, dl1b, db1 = na_bwd(d_a1, h1, l1w, l1b, mean2, rstd2, skip_shape=(rows, E), eps=eps[1], act=1,
                                             drop_p=p1, seed=seed1, colsum=True, d_out_pool=d_a1p, pool_k=pk)
        dw1 = torch.mm(d_h1.t(), x1)
        d_x1.addmm_(d_h1, w1c)
        # x1 = LN1(m) + x: upstream = d_x1 (skip + GEMM paths) + d_z (LN2 path), summed inside the kernel;
        # the identity skip hands the summed gradient on as d_x
        d_m, d_x, dn1w, dn1b, _ = na_bwd(d_x1, m.view(rows, E), n1w, n1b, mean1, rstd1, skip_shape=(rows, E), eps=eps[0],
                                         dout2=d_z)
        d_xm, d_mp = (fftmix_bwd_raw if kind == "fft" else dwtmix_bwd_raw)(d_m.view(B, T, E), x, mp, H, W, has_cls)
        d_x = d_x.view(B, T, E)
        if ctx.has_twin:
            d_twin = d_xm                   # the layer below adds the two parts inside its LN2 backward
        else:
            d_x, d_twin = d_x.add_(d_xm), None
        wd1, bd1, wd3, bd3 = ctx.param_dtypes
        return (d_x, d_twin, d_mp, dn1w, dn1b, dw1.to(wd1), db1.to(bd1), dl1w, dl1b, dw3.to(wd3), db3.to(bd3), dl3w, dl3b,
                dn2w, dn2b, None)


def encoder_layer(x, mix_kind, mix_param, grid, has_cls, norm1, lin1, ln1, lin3, ln3, norm2, p1, p2, seed1, seed2,
                  x_twin=None, return_twin=False):
    """Whole ``SpectreEncoderLayer`` through the fused kernels as one autograd node.  ``x`` is ``[B, T, E]`` in the
    stream dtype (float32 or bfloat16); modules are the layer's ``nn.LayerNorm`` / ``nn.Linear`` instances.

    Stacked layers pass ``(y, y_twin)`` of one layer as ``(x, x_twin)`` of the next (``return_twin=True``): the twin
    is a second view of the same values that only carries the second addend of the input gradient backwards."""
    cfg = (mix_kind, int(grid[0]), int(grid[1]), bool(has_cls), (norm1.eps, ln1.eps, ln3.eps, norm2.eps),
           float(p1), float(p2), int(seed1), int(seed2))
    y, twin = _EncoderLayer.apply(x, x_twin, mix_param, norm1.weight, norm1.bias, lin1.weight, lin1.bias, ln1.weight, ln1.bias,
                               lin3.weight, lin3.bias, ln3.weight, ln3.bias, norm2.weight, norm2.bias, cfg)
    return (y, twin) if return_twin else y
