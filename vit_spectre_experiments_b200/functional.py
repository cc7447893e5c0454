"""Autograd functions over the C-ABI kernels (``include/spectre_b200.h``).

Forward and backward of the occupant of the reference's ``mix_layer`` slot
(``spectre_vit/models/spectre/spectre.py:50,66``).  torch is plumbing only here: it owns the
device buffers and the current stream; the arithmetic happens in ``libspectre_b200.so``.
"""
from __future__ import annotations

import torch

from . import _lib

_DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1}


class KernelTimer:
    """Optional CUDA-event bracketing of every C-ABI call on the launching stream (bench.py uses it to
    time the mixer kernels live inside the training step).  Off by default: zero overhead."""

    def __init__(self):
        self.enabled = False
        self.records = {}   # name -> list of (start_event, end_event, meta, main_kernel_end_event)

    def reset(self):
        self.records = {}

    def summary(self):
        """name -> (calls, mean ms of the whole C-ABI call, meta of the last call, mean ms of the call's main
        kernel alone); synchronises.  The last value differs from the second only for backward mixer calls, which
        end with a few-microsecond reduction kernel of the per-slab parameter gradients."""
        torch.cuda.synchronize()
        out = {}
        for name, evs in self.records.items():
            ms = [s.elapsed_time(e) for s, e, _, _ in evs]
            main = [s.elapsed_time(m) for s, _, _, m in evs]
            n = max(len(ms), 1)
            out[name] = (len(ms), sum(ms) / n, evs[-1][2], sum(main) / n)
        return out


TIMER = KernelTimer()


class _timed:
    def __init__(self, name, meta, mark_main=False):
        self.name, self.meta, self.mark_main = name, meta, mark_main

    def __enter__(self):
        if TIMER.enabled:
            self.s = torch.cuda.Event(enable_timing=True)
            self.e = torch.cuda.Event(enable_timing=True)
            self.s.record()
            if self.mark_main:   # the library re-records this event between the main kernel and the reduction
                self.m = torch.cuda.Event(enable_timing=True)
                self.m.record()  # creates the underlying cudaEvent_t
                _lib.load().spectre_debug_mark_main_kernel_end(self.m.cuda_event)
        return self

    def __exit__(self, *exc):
        if TIMER.enabled:
            self.e.record()
            main_end = self.e
            if self.mark_main and not _lib.load().spectre_debug_mark_pending():
                main_end = self.m
            TIMER.records.setdefault(self.name, []).append((self.s, self.e, self.meta, main_end))
        return False


def _act_code(t: torch.Tensor, what: str) -> int:
    if not t.is_cuda:
        raise RuntimeError(
            f"{what}: tensor is on {t.device}; the spectral mixers are CUDA (sm_100a) only and have no CPU fallback"
        )
    try:
        return _DTYPE_CODE[t.dtype]
    except KeyError:
        raise RuntimeError(f"{what}: activations must be float32 or bfloat16, got {t.dtype}") from None


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _check_tokens(x: torch.Tensor, H: int, W: int, has_cls: bool, what: str) -> None:
    if x.dim() != 3 or x.shape[1] != H * W + int(has_cls):
        raise RuntimeError(
            f"{what}: expected x of shape [B, {H * W + int(has_cls)}, E] for a {H}x{W} grid"
            f"{' + CLS' if has_cls else ''}, got {tuple(x.shape)}"
        )


# ------------------------------------------------------------------ raw (non-autograd) kernel calls
def _mix_fwd_raw(kind, x, param, H, W, has_cls):
    lib = _lib.load()
    what = f"{kind}_mix"
    _check_tokens(x, H, W, has_cls, what)
    code = _act_code(x, what)
    E = x.shape[2]
    want = (H, W // 2 + 1, E, 2) if kind == "fft" else (4, (H + 1) // 2, (W + 1) // 2, E)
    if param.dtype != torch.float32 or tuple(param.shape) != want:
        name = "filter" if kind == "fft" else "weight"
        raise RuntimeError(f"{what}: {name} must be float32 {list(want)}, got {param.dtype} {tuple(param.shape)}")
    y = torch.empty_like(x)
    B = x.shape[0]
    fn = lib.spectre_fftmix_fwd if kind == "fft" else lib.spectre_dwtmix_fwd
    with torch.cuda.device(x.device):
        with _timed(f"{kind}mix_fwd", (B, H, W, E, x.element_size())):
            rc = fn(x.data_ptr(), param.data_ptr(), y.data_ptr(), B, H, W, E, int(has_cls), code, _stream_ptr(x.device))
    _lib.check(rc, f"spectre_{kind}mix_fwd")
    _lib.count_launches()
    return y


def _mix_bwd_raw(kind, dy, x, param, H, W, has_cls):
    lib = _lib.load()
    B, _, E = x.shape
    dx = torch.empty_like(x)
    dparam = torch.empty_like(param)
    ws_fn = lib.spectre_fftmix_bwd_workspace_bytes if kind == "fft" else lib.spectre_dwtmix_bwd_workspace_bytes
    fn = lib.spectre_fftmix_bwd if kind == "fft" else lib.spectre_dwtmix_bwd
    with torch.cuda.device(x.device):
        nbytes = ws_fn(B, H, W, E)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
        with _timed(f"{kind}mix_bwd", (B, H, W, E, x.element_size()), mark_main=True):
            rc = fn(dy.data_ptr(), x.data_ptr(), param.data_ptr(), dx.data_ptr(), dparam.data_ptr(), ws.data_ptr(),
                    nbytes, B, H, W, E, int(has_cls), _act_code(x, f"{kind}_mix"), _stream_ptr(x.device))
    _lib.check(rc, f"spectre_{kind}mix_bwd")
    _lib.count_launches()
    return dx, dparam


def fftmix_fwd_raw(x, filt, H, W, has_cls):
    """``spectre_fftmix_fwd`` on contiguous ``x [B,T,E]`` (fp32 / bf16) and float32 ``filt``; returns ``y``."""
    return _mix_fwd_raw("fft", x, filt, H, W, has_cls)


def fftmix_bwd_raw(dy, x, filt, H, W, has_cls):
    """``spectre_fftmix_bwd``: returns ``(dx, dfilt)``."""
    return _mix_bwd_raw("fft", dy, x, filt, H, W, has_cls)


def dwtmix_fwd_raw(x, weight, H, W, has_cls):
    return _mix_fwd_raw("dwt", x, weight, H, W, has_cls)


def dwtmix_bwd_raw(dy, x, weight, H, W, has_cls):
    return _mix_bwd_raw("dwt", dy, x, weight, H, W, has_cls)


class _Mix(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, param, kind, H, W, has_cls):
        x = x.contiguous()
        param = param.contiguous()
        y = _mix_fwd_raw(kind, x, param, H, W, has_cls)
        ctx.save_for_backward(x, param)
        ctx.cfg = (kind, H, W, has_cls)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, param = ctx.saved_tensors
        kind, H, W, has_cls = ctx.cfg
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dx, dparam = _mix_bwd_raw(kind, dy, x, param, H, W, has_cls)
        return dx, dparam, None, None, None, None


class _PermutMix(torch.autograd.Function):
    """``x.view(B, -1)[:, perms] * signs`` (layers.py:68-70) through ``spectre_permut_fwd`` / ``_bwd``."""

    @staticmethod
    def forward(ctx, x, perms, signs, inv):
        lib = _lib.load()
        code = _act_code(x, "permut_mix")
        x = x.contiguous()
        B = x.shape[0]
        heads, F = perms.shape
        if x.numel() != B * F:
            raise RuntimeError(f"permut_mix: x has {x.numel() // B} elements per image, perms index {F}")
        if perms.dtype != torch.int64 or signs.dtype != torch.float32 or signs.numel() != heads * F:
            raise RuntimeError("permut_mix: perms must be int64 [heads, T*E], signs float32 [heads, T*E]")
        out = torch.empty(B, heads, F, dtype=x.dtype, device=x.device)
        with torch.cuda.device(x.device):
            with _timed("permut_fwd", (B, heads, F, x.element_size())):
                rc = lib.spectre_permut_fwd(x.data_ptr(), perms.data_ptr(), signs.data_ptr(), out.data_ptr(), B, F, heads,
                                            code, _stream_ptr(x.device))
        _lib.check(rc, "spectre_permut_fwd")
        _lib.count_launches()
        ctx.save_for_backward(inv, signs)
        ctx.shape = x.shape
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.load()
        inv, signs = ctx.saved_tensors
        heads, F = inv.shape
        dout = dout.contiguous()
        B = dout.shape[0]
        dx = torch.empty(ctx.shape, dtype=dout.dtype, device=dout.device)
        with torch.cuda.device(dout.device):
            with _timed("permut_bwd", (B, heads, F, dout.element_size())):
                rc = lib.spectre_permut_bwd(dout.data_ptr(), inv.data_ptr(), signs.data_ptr(), dx.data_ptr(), B, F, heads,
                                            _act_code(dout, "permut_mix"), _stream_ptr(dout.device))
        _lib.check(rc, "spectre_permut_bwd")
        _lib.count_launches()
        return dx, None, None, None


def permut_mix(x: torch.Tensor, perms: torch.Tensor, signs: torch.Tensor, inv_perms: torch.Tensor) -> torch.Tensor:
    """``(x.view(B, -1)[:, perms] * signs)`` -> ``[B, heads, T*E]`` in x's dtype (the sign flip is exact).

    perms int64 ``[heads, T*E]`` with permutation rows, signs float32 ``[heads, T*E]`` (or ``[1, heads, T*E]``),
    inv_perms the row-wise inverse permutations (``inverse_permutations(perms)``)."""
    return _PermutMix.apply(x, perms.contiguous(), signs.reshape(perms.shape).contiguous(), inv_perms.contiguous())


def inverse_permutations(perms: torch.Tensor) -> torch.Tensor:
    """Row-wise inverse of ``perms [heads, F]``; raises if a row is not a permutation of ``range(F)``."""
    heads, F = perms.shape
    ar = torch.arange(F, device=perms.device, dtype=perms.dtype).expand(heads, F)
    inv = torch.full_like(perms, -1)
    inv.scatter_(1, perms, ar)
    if bool((inv < 0).any()) or int(perms.min()) < 0 or int(perms.max()) >= F:
        raise RuntimeError("permut_mix: every row of perms must be a permutation of range(T*E)")
    return inv


class _PatchEmbed(torch.autograd.Function):
    """``dropout(cat(cls, patches @ M + b) + pos)`` (spectre.py:124-156 with the DFT / gains / projection folded into
    ``M``): patch gather + cast, the library GEMM, and the cls / position / dropout pass -- three launches forward,
    four backward (``spectre_patch_gather``, ``spectre_embed_finish_fwd/bwd``)."""

    @staticmethod
    def forward(ctx, img, M, bias, cls, pos, P, drop_p, seed, dt):
        lib = _lib.load()
        if not img.is_cuda:
            raise RuntimeError("patch_embed: CUDA (sm_100a) only; there is no CPU fallback")
        img = img.float().contiguous()
        B, C, Hi, Wi = img.shape
        nH, nW = Hi // P, Wi // P
        n, K, E = nH * nW, C * P * P, M.shape[1]
        code = _DTYPE_CODE[dt]
        A = torch.empty(B * n, K, dtype=dt, device=img.device)
        st = _stream_ptr(img.device)
        with torch.cuda.device(img.device):
            with _timed("patch_gather", (B * n, K, A.element_size(), img.numel() * 4 + A.numel() * A.element_size())):
                rc = lib.spectre_patch_gather(img.data_ptr(), A.data_ptr(), B, C, Hi, Wi, P, code, st)
            _lib.check(rc, "spectre_patch_gather")
            _lib.count_launches()
            Md = M.detach().to(dt)
            tok = torch.addmm(bias.detach().to(dt), A, Md)
            out = torch.empty(B, n + 1, E, dtype=dt, device=img.device)
            clsf, posf = cls.detach().float().contiguous().view(-1), pos.detach().float().contiguous().view(n + 1, E)
            with _timed("embed_finish_fwd", (B * (n + 1), E, out.element_size(), (tok.numel() + out.numel()) * out.element_size())):
                rc = lib.spectre_embed_finish_fwd(tok.data_ptr(), clsf.data_ptr(), posf.data_ptr(), out.data_ptr(), B, n + 1, E,
                                                  code, float(drop_p), int(seed), _lib.seed_dev_ptr() if drop_p > 0 else None, st)
            _lib.check(rc, "spectre_embed_finish_fwd")
            _lib.count_launches()
        ctx.save_for_backward(A, Md)
        ctx.meta = (B, n, E, float(drop_p), int(seed), code, M.dtype, bias.dtype, cls.shape, pos.shape, cls.dtype, pos.dtype)
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.load()
        A, Md = ctx.saved_tensors
        B, n, E, drop_p, seed, code, m_dt, b_dt, cls_shape, pos_shape, cls_dt, pos_dt = ctx.meta
        dout = dout.to(A.dtype).contiguous()
        dtok = torch.empty(B * n, E, dtype=A.dtype, device=A.device)
        dpos = torch.empty(n + 1, E, dtype=torch.float32, device=A.device)
        with torch.cuda.device(A.device):
            nws = lib.spectre_embed_finish_bwd_workspace_bytes(B, n + 1, E)
            ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=A.device)
            with _timed("embed_finish_bwd", (B * (n + 1), E, dout.element_size(), (dout.numel() + dtok.numel()) * dout.element_size())):
                rc = lib.spectre_embed_finish_bwd(dout.data_ptr(), dtok.data_ptr(), dpos.data_ptr(), ws.data_ptr(), nws, B, n + 1, E,
                                                  code, drop_p, seed, _lib.seed_dev_ptr() if drop_p > 0 else None,
                                                  _stream_ptr(A.device))
            _lib.check(rc, "spectre_embed_finish_bwd")
            _lib.count_launches()
        dM = torch.mm(A.t(), dtok).to(m_dt)
        dbias = dpos[1:].sum(0).to(b_dt)
        return (None, dM, dbias, dpos[0].reshape(cls_shape).to(cls_dt), dpos.reshape(pos_shape).to(pos_dt),
                None, None, None, None)


def patch_embed(img, M, bias, cls, pos, patch: int, drop_p: float = 0.0, seed: int = 0, dtype=torch.float32) -> torch.Tensor:
    """``dropout(cat(cls, patches(img) @ M + bias) + pos)`` -> ``[B, 1 + n, E]`` in ``dtype`` (float32 / bfloat16).

    img float32 ``[B, C, Hi, Wi]``; ``M [C*P*P, E]`` the folded Re(rfft2) x gains x projection matrix (rows ordered
    (c, i, j)); cls ``[1, 1, E]``, pos ``[1, 1 + n, E]``.  Gradients flow to ``M``, ``bias``, ``cls`` and ``pos``."""
    return _PatchEmbed.apply(img, M, bias, cls, pos, int(patch), float(drop_p), int(seed), dtype)


def fft_mix(x: torch.Tensor, filt: torch.Tensor, grid, has_cls: bool = True) -> torch.Tensor:
    """``cat(cls, irfft2(rfft2(x_grid, ortho) * filt, ortho))`` in one fused CUDA pass.

    x ``[B, has_cls + H*W, E]`` (float32 or bfloat16, CUDA); filt float32 ``[H, W//2+1, E, 2]``.
    """
    H, W = grid
    return _Mix.apply(x, filt, "fft", int(H), int(W), bool(has_cls))


def dwt_mix(x: torch.Tensor, weight: torch.Tensor, grid, has_cls: bool = True) -> torch.Tensor:
    """``cat(cls, IDWT_haar(weight * DWT_haar(x_grid)))`` in one fused CUDA pass.

    weight float32 ``[4, ceil(H/2), ceil(W/2), E]`` ordered (LL, W-low/H-high, W-high/H-low, HH); odd grids are
    zero-padded (``mode="zero"``) and cropped.
    """
    H, W = grid
    return _Mix.apply(x, weight, "dwt", int(H), int(W), bool(has_cls))


class _DwtMl(torch.autograd.Function):
    """Multi-level Haar mixer (``spectre_dwtmix_ml_fwd/bwd``); weight float32 ``[H, W, E]`` in the in-place layout."""

    @staticmethod
    def forward(ctx, x, weight, H, W, has_cls, levels):
        lib = _lib.load()
        what = "dwt_mix_ml"
        _check_tokens(x, H, W, has_cls, what)
        code = _act_code(x, what)
        B, _, E = x.shape
        if weight.dtype != torch.float32 or tuple(weight.shape) != (H, W, E):
            raise RuntimeError(f"{what}: weight must be float32 {[H, W, E]}, got {weight.dtype} {tuple(weight.shape)}")
        x = x.contiguous()
        w = weight.detach().contiguous()
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            with _timed("dwtmlmix_fwd", (B, H, W, E, x.element_size())):
                rc = lib.spectre_dwtmix_ml_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, H, W, E, int(has_cls),
                                               int(levels), code, _stream_ptr(x.device))
        _lib.check(rc, "spectre_dwtmix_ml_fwd")
        _lib.count_launches()
        ctx.save_for_backward(x, w)
        ctx.meta = (H, W, has_cls, levels)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, w = ctx.saved_tensors
        H, W, has_cls, levels = ctx.meta
        B, _, E = x.shape
        dy = dy.to(x.dtype).contiguous()
        dx, dw = torch.empty_like(x), torch.empty_like(w)
        with torch.cuda.device(x.device):
            nbytes = lib.spectre_dwtmix_ml_bwd_workspace_bytes(B, H, W, E)
            ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
            with _timed("dwtmlmix_bwd", (B, H, W, E, x.element_size()), mark_main=True):
                rc = lib.spectre_dwtmix_ml_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                                               ws.data_ptr(), nbytes, B, H, W, E, int(has_cls), int(levels),
                                               _act_code(x, "dwt_mix_ml"), _stream_ptr(x.device))
        _lib.check(rc, "spectre_dwtmix_ml_bwd")
        _lib.count_launches()
        return dx, dw, None, None, None, None


def dwt_mix_ml(x: torch.Tensor, weight: torch.Tensor, grid, has_cls: bool = True, levels: int = 3) -> torch.Tensor:
    """``cat(cls, T^-1(weight * T(x_grid)))`` with ``T`` the ``levels``-level orthonormal 2-D Haar transform
    (``DWTForward(J=3, wave="haar", mode="zero")``, dwt_experiments.py:56), one fused CUDA pass.

    weight float32 ``[H, W, E]``: coefficients addressed in place on the token lattice (include/spectre_b200.h).
    """
    H, W = grid
    return _DwtMl.apply(x, weight, int(H), int(W), bool(has_cls), int(levels))


_HAAR1D_AXIS = {"token": 0, "embed": 1}


def haar1d_weight_shape(n_tokens: int, embed_dim: int, axis: str):
    """Shape of the float32 weight of ``haar1d_mix``: (low, high) sub-bands of the 1-D Haar along ``axis``."""
    return (2, (n_tokens + 1) // 2, embed_dim) if axis == "token" else (2, n_tokens, embed_dim // 2)


class _Haar1d(torch.autograd.Function):
    """1-D Haar mixers ``dwt_token`` / ``dwt_embed`` (``spectre_haar1d_fwd/bwd``; mode names spectre.py:33-34)."""

    @staticmethod
    def forward(ctx, x, weight, has_cls, axis):
        lib = _lib.load()
        what = f"haar1d_mix({axis})"
        if axis not in _HAAR1D_AXIS:
            raise RuntimeError(f"{what}: axis must be 'token' or 'embed'")
        code = _act_code(x, what)
        if x.dim() != 3 or x.shape[1] <= int(has_cls):
            raise RuntimeError(f"{what}: expected x [B, {'1 + ' if has_cls else ''}N, E], got {tuple(x.shape)}")
        B, T, E = x.shape
        N = T - int(has_cls)
        if E % 8:
            raise RuntimeError(f"{what}: embed_dim must be a multiple of 8, got {E}")
        want = haar1d_weight_shape(N, E, axis)
        if weight.dtype != torch.float32 or tuple(weight.shape) != want:
            raise RuntimeError(f"{what}: weight must be float32 {list(want)}, got {weight.dtype} {tuple(weight.shape)}")
        x = x.contiguous()
        w = weight.detach().contiguous()
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            with _timed("haar1d_fwd", (B, N, 1, E, x.element_size())):
                rc = lib.spectre_haar1d_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, N, E, int(has_cls),
                                            _HAAR1D_AXIS[axis], code, _stream_ptr(x.device))
        _lib.check(rc, "spectre_haar1d_fwd")
        _lib.count_launches()
        ctx.save_for_backward(x, w)
        ctx.meta = (N, has_cls, axis)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, w = ctx.saved_tensors
        N, has_cls, axis = ctx.meta
        B, _, E = x.shape
        dy = dy.to(x.dtype).contiguous()
        dx, dw = torch.empty_like(x), torch.empty_like(w)
        with torch.cuda.device(x.device):
            nbytes = lib.spectre_haar1d_bwd_workspace_bytes(B, N, E, _HAAR1D_AXIS[axis])
            ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
            with _timed("haar1d_bwd", (B, N, 1, E, x.element_size())):
                rc = lib.spectre_haar1d_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                                            ws.data_ptr(), nbytes, B, N, E, int(has_cls), _HAAR1D_AXIS[axis],
                                            _act_code(x, "haar1d_mix"), _stream_ptr(x.device))
        _lib.check(rc, "spectre_haar1d_bwd")
        _lib.count_launches()
        return dx, dw, None, None


def haar1d_mix(x: torch.Tensor, weight: torch.Tensor, has_cls: bool = True, axis: str = "token") -> torch.Tensor:
    """``cat(cls, IHaar(weight * Haar(x[:, cls:])))`` with the single-level orthonormal 1-D Haar along the token axis
    (``axis="token"``, the reference's ``dwt_token`` mode name) or the embedding axis (``"embed"``, ``dwt_embed``),
    spectre.py:33-34; one fused CUDA pass.  weight float32 ``haar1d_weight_shape(N, E, axis)``; an odd token count is
    zero-padded (``mode="zero"``) and cropped."""
    return _Haar1d.apply(x, weight, bool(has_cls), str(axis))
