"""Oracle (CPU restatement) of the reference's spectral token-mixer path.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.

What it restates
----------------
The reference (Biblbrox/ViT-Spectre-Experiments @ fa0d82be) has no module that
composes ``rfft2 -> learned complex filter -> irfft2``;
the "reference path" is the composition of the reference's own primitives and
conventions placed in its ``mix_layer`` slot:

* slot + call:   ``spectre_vit/models/spectre/spectre.py:50`` (construction),
                 ``:66`` (``x = self.norm1(self.mix_layer(x)) + x``);
                 tokens are ``[B, 1 + H*W, E]`` with CLS first (``:151-152,182,198``)
* 2-D FFT conv.: ``torch.fft.rfft2(x, norm="ortho")`` at ``spectre.py:136``;
                 2-D FFT over the token grid as the mixer at
                 ``spectre_vit/repl/orthogonal_permut.py:23-28``
* 1-D FFT conv.: ``torch.fft.rfft(x, dim=-1).real`` at ``spectre_vit/modules/spectre.py:14``
* Haar / zero:   ``DWTForward(J=3, wave="haar", mode="zero")`` at
                 ``spectre_vit/repl/dwt_experiments.py:9,56``

Pinning status
--------------
* FFT mixer: PINNED.  ``torch.fft`` (the reference's arithmetic) is executed
  directly, and an independent float64 direct-DFT restatement
  (:func:`rfft2_ortho_np`, :func:`fft_mix_np`) is checked against (a) torch.fft
  and (b) golden vectors produced by importing the reference's own
  ``SpectralPatchEmbed`` / ``FFT`` / ``SpectreViT`` classes
  (``oracle/make_golden.py`` -> ``tests/golden/*.npz``).
* Haar DWT mixer: PARITY UNPINNED.  The arithmetic would live in
  ``pytorch-wavelets>=1.3.0`` / ``ptwt>=1.0.1`` / ``pywavelets>=1.9.0``
  (``pyproject.toml:11,23,27,28``), none vendored under the reference, none
  installed here, no version pinned (uv.lock is git-ignored) and the reference
  holds no test or fixture at that boundary.  The restatement follows the
  published orthonormal Haar analysis filters ``lo=[1,1]/sqrt2``,
  ``hi=[1,-1]/sqrt2`` and is checked against an explicit conv filter bank and
  perfect reconstruction only.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as nn

# --------------------------------------------------------------------------
# FFT mixer
# --------------------------------------------------------------------------


def as_complex_filter(filt: torch.Tensor) -> torch.Tensor:
    """``[H, Wh, E, 2]`` real parameter -> ``[H, Wh, E]`` complex view."""
    if torch.is_complex(filt):
        return filt
    return torch.view_as_complex(filt.contiguous())


def fft_mix_ref(x: torch.Tensor, filt: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """y = cat(cls, irfft2(rfft2(grid) * filt)) with the reference's ortho norm.

    x: ``[B, 1+H*W, E]`` fp32/fp64; filt: complex ``[H, W//2+1, E]`` or real
    ``[H, W//2+1, E, 2]``.  Conventions: ``spectre.py:136`` (``norm="ortho"``),
    grid = tokens 1.. reshaped ``[B, H, W, E]`` (``spectre.py:151-152``).
    """
    B, T, E = x.shape
    assert T == 1 + H * W
    f = as_complex_filter(filt)
    g = x[:, 1:].reshape(B, H, W, E)
    X = torch.fft.rfft2(g, dim=(1, 2), norm="ortho")
    y = torch.fft.irfft2(X * f, s=(H, W), dim=(1, 2), norm="ortho")
    return torch.cat([x[:, :1], y.reshape(B, H * W, E)], dim=1)


def fft_mix_bwd_ref(dy: torch.Tensor, x: torch.Tensor, filt: torch.Tensor, H: int, W: int):
    """Closed-form backward of :func:`fft_mix_ref`.

    Returns ``(dx [B,T,E], dfilt [H,Wh,E,2])``.  ``dfilt`` is the gradient w.r.t.
    the real ``[H,Wh,E,2]`` parameter, i.e. (Re, Im) of
    ``sum_b conj(X_b) * (c * DY_b)`` with ``c[kw]=1`` for kw=0 and (W even)
    kw=W/2, else 2.
    """
    B, T, E = x.shape
    f = as_complex_filter(filt)
    Wh = W // 2 + 1
    DY = torch.fft.rfft2(dy[:, 1:].reshape(B, H, W, E), dim=(1, 2), norm="ortho")
    X = torch.fft.rfft2(x[:, 1:].reshape(B, H, W, E), dim=(1, 2), norm="ortho")
    dxg = torch.fft.irfft2(f.conj() * DY, s=(H, W), dim=(1, 2), norm="ortho")
    c = torch.full((Wh,), 2.0, dtype=x.dtype, device=x.device)
    c[0] = 1.0
    if W % 2 == 0:
        c[-1] = 1.0
    dfc = (X.conj() * DY * c.view(1, 1, Wh, 1)).sum(0)
    dx = torch.cat([dy[:, :1], dxg.reshape(B, H * W, E)], dim=1)
    return dx, torch.view_as_real(dfc).contiguous()


def rfft2_ortho_np(g: np.ndarray) -> np.ndarray:
    """Independent float64 direct-DFT restatement of
    ``torch.fft.rfft2(g, dim=(-2,-1), norm="ortho")`` (no FFT library)."""
    g = np.asarray(g, dtype=np.float64)
    H, W = g.shape[-2:]
    Wh = W // 2 + 1
    Fw = np.exp(-2j * np.pi * np.outer(np.arange(W), np.arange(Wh)) / W)      # [W, Wh]
    Fh = np.exp(-2j * np.pi * np.outer(np.arange(H), np.arange(H)) / H)       # [kh, h]
    rows = g @ Fw                                                             # [..., H, Wh]
    return np.einsum("kh,...hw->...kw", Fh, rows) / math.sqrt(H * W)


def irfft2_ortho_np(Z: np.ndarray, H: int, W: int) -> np.ndarray:
    """Direct-DFT restatement of ``torch.fft.irfft2(Z, s=(H,W), norm="ortho")``
    for a possibly non-Hermitian ``Z`` ``[..., H, Wh]``: complex inverse along
    H, then C2R along W which ignores Im of column 0 and (W even) column W/2."""
    Z = np.asarray(Z, dtype=np.complex128)
    Wh = W // 2 + 1
    Fh = np.exp(2j * np.pi * np.outer(np.arange(H), np.arange(H)) / H)        # [h, kh]
    cols = np.einsum("hk,...kw->...hw", Fh, Z)                                # [..., H, Wh]
    c = np.full(Wh, 2.0)
    c[0] = 1.0
    if W % 2 == 0:
        c[-1] = 1.0
    cols = cols.copy()
    cols[..., 0] = cols[..., 0].real
    if W % 2 == 0:
        cols[..., -1] = cols[..., -1].real
    Fw = np.exp(2j * np.pi * np.outer(np.arange(Wh), np.arange(W)) / W)       # [kw, w]
    return ((cols * c) @ Fw).real / math.sqrt(H * W)


def fft_mix_np(x: np.ndarray, filt: np.ndarray, H: int, W: int) -> np.ndarray:
    """Float64 direct-DFT version of :func:`fft_mix_ref`.  ``filt`` complex ``[H,Wh,E]``."""
    x = np.asarray(x, dtype=np.float64)
    B, T, E = x.shape
    g = x[:, 1:].reshape(B, H, W, E).transpose(0, 3, 1, 2)                    # [B,E,H,W]
    X = rfft2_ortho_np(g)
    Z = X * np.asarray(filt).transpose(2, 0, 1)[None]
    y = irfft2_ortho_np(Z, H, W).transpose(0, 2, 3, 1).reshape(B, H * W, E)
    return np.concatenate([x[:, :1], y], axis=1)


# --------------------------------------------------------------------------
# Haar DWT mixer (rows A5 / A6) -- parity unpinned, see module docstring
# --------------------------------------------------------------------------


def haar2d(g: torch.Tensor) -> torch.Tensor:
    """Single-level orthonormal 2-D Haar analysis.  g ``[B,H,W,E]`` (H, W even)
    -> ``[B,4,H/2,W/2,E]`` ordered (LL, W-low/H-high, W-high/H-low, HH)."""
    a, b = g[:, 0::2, 0::2], g[:, 0::2, 1::2]
    c, d = g[:, 1::2, 0::2], g[:, 1::2, 1::2]
    return torch.stack(
        [(a + b + c + d) / 2, (a + b - c - d) / 2, (a - b + c - d) / 2, (a - b - c + d) / 2], dim=1
    )


def ihaar2d(s4: torch.Tensor) -> torch.Tensor:
    """Inverse of :func:`haar2d` (same butterflies; the 4x4 block matrix is
    orthogonal and symmetric)."""
    ll, s1, s2, hh = s4.unbind(1)
    a = (ll + s1 + s2 + hh) / 2
    b = (ll + s1 - s2 - hh) / 2
    c = (ll - s1 + s2 - hh) / 2
    d = (ll - s1 - s2 + hh) / 2
    B, h, w, E = a.shape
    out = a.new_empty(B, 2 * h, 2 * w, E)
    out[:, 0::2, 0::2], out[:, 0::2, 1::2] = a, b
    out[:, 1::2, 0::2], out[:, 1::2, 1::2] = c, d
    return out


def _pad_even(g: torch.Tensor) -> torch.Tensor:
    """``mode="zero"`` for a length-2 filter: an odd axis gets one zero appended at its end
    (pytorch_wavelets pads ``p = 2*ceil(N/2) - N`` samples, ``p // 2`` in front and the rest behind)."""
    B, H, W, E = g.shape
    if H % 2 == 0 and W % 2 == 0:
        return g
    out = g.new_zeros(B, H + H % 2, W + W % 2, E)
    out[:, :H, :W] = g
    return out


def dwt_mix_ref(x: torch.Tensor, w: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """y = cat(cls, crop(IDWT(w * DWT(pad(grid))))).  w real ``[4, ceil(H/2), ceil(W/2), E]``."""
    B, T, E = x.shape
    assert T == 1 + H * W
    y = ihaar2d(haar2d(_pad_even(x[:, 1:].reshape(B, H, W, E))) * w)[:, :H, :W]
    return torch.cat([x[:, :1], y.reshape(B, H * W, E)], dim=1)


def dwt_mix_bwd_ref(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, H: int, W: int):
    """Closed-form backward (row A6): dx = IDWT(w * DWT(dy)); dw = sum_b DWT(x) * DWT(dy)."""
    B, T, E = x.shape
    DY = haar2d(_pad_even(dy[:, 1:].reshape(B, H, W, E)))
    X = haar2d(_pad_even(x[:, 1:].reshape(B, H, W, E)))
    dxg = ihaar2d(DY * w)[:, :H, :W]       # crop^T = pad, and the block transform is orthogonal and symmetric
    dx = torch.cat([dy[:, :1], dxg.reshape(B, H * W, E)], dim=1)
    return dx, (X * DY).sum(0)


def haar_ml_inplace(g: torch.Tensor, levels: int, inverse: bool = False) -> torch.Tensor:
    """J-level orthonormal 2-D Haar with the coefficients left IN PLACE on the lattice (``DWTForward(J=3,
    wave="haar", mode="zero")``, dwt_experiments.py:56; H, W multiples of 2^J so no padding arises).  Level l works on
    the 2x2 blocks (a b / c d) of the points whose coordinates are multiples of 2^(l-1): LL -> a, low-W/high-H -> c,
    high-W/low-H -> b, HH -> d; the next level recurses on the LL points.  ``inverse`` undoes it (levels J..1)."""
    out = g.clone()
    order = range(levels, 0, -1) if inverse else range(1, levels + 1)
    for l in order:
        st, h = 1 << l, 1 << (l - 1)
        a, b = out[:, 0::st, 0::st].clone(), out[:, 0::st, h::st].clone()
        c, d = out[:, h::st, 0::st].clone(), out[:, h::st, h::st].clone()
        if inverse:                      # stored (ll, s2, s1, hh) at (a, b, c, d)
            ll, s2, s1, hh = a, b, c, d
            out[:, 0::st, 0::st] = (ll + s1 + s2 + hh) / 2
            out[:, 0::st, h::st] = (ll + s1 - s2 - hh) / 2
            out[:, h::st, 0::st] = (ll - s1 + s2 - hh) / 2
            out[:, h::st, h::st] = (ll - s1 - s2 + hh) / 2
        else:
            out[:, 0::st, 0::st] = (a + b + c + d) / 2
            out[:, h::st, 0::st] = (a + b - c - d) / 2     # low along W, high along H
            out[:, 0::st, h::st] = (a - b + c - d) / 2     # high along W, low along H
            out[:, h::st, h::st] = (a - b - c + d) / 2
    return out


def dwt_ml_mix_ref(x: torch.Tensor, w: torch.Tensor, H: int, W: int, levels: int) -> torch.Tensor:
    """y = cat(cls, T^-1(w * T(grid))), T = ``levels``-level Haar; w real ``[H, W, E]`` (in-place layout)."""
    B, T, E = x.shape
    assert T == 1 + H * W and H % (1 << levels) == 0 and W % (1 << levels) == 0
    y = haar_ml_inplace(haar_ml_inplace(x[:, 1:].reshape(B, H, W, E), levels) * w, levels, inverse=True)
    return torch.cat([x[:, :1], y.reshape(B, H * W, E)], dim=1)


def dwt_ml_mix_bwd_ref(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, H: int, W: int, levels: int):
    """dx = T^-1(w * T dy) (T orthogonal); dw = sum_b T(x_b) * T(dy_b)."""
    B, T, E = x.shape
    DY = haar_ml_inplace(dy[:, 1:].reshape(B, H, W, E), levels)
    X = haar_ml_inplace(x[:, 1:].reshape(B, H, W, E), levels)
    dxg = haar_ml_inplace(DY * w, levels, inverse=True)
    return torch.cat([dy[:, :1], dxg.reshape(B, H * W, E)], dim=1), (X * DY).sum(0)


# ------------------------------------------------------------------ 1-D Haar mixers ("dwt_token" / "dwt_embed")
# spectre_vit/models/spectre/spectre.py:33-34 names the two modes ("Wavelet transform in token dimension" / "in embedding
# dimension") and the reference has no code for them: the operator is the 1-D analogue of dwt_mix_ref above -- parity
# unpinned, like every Haar function of this file.
def _haar1d_pairs(g: torch.Tensor, axis: str):
    """(low, high) of the single-level orthonormal Haar along the token axis (odd count: a zero token appended,
    mode="zero") or the embedding axis of ``g [B, N, E]``."""
    r = 2.0 ** -0.5
    if axis == "token":
        gp = torch.nn.functional.pad(g, (0, 0, 0, g.shape[1] % 2))
        a, b = gp[:, 0::2], gp[:, 1::2]
    else:
        a, b = g[..., 0::2], g[..., 1::2]
    return (a + b) * r, (a - b) * r


def _ihaar1d_pairs(lo: torch.Tensor, hi: torch.Tensor, axis: str, n_tokens: int) -> torch.Tensor:
    r = 2.0 ** -0.5
    a, b = (lo + hi) * r, (lo - hi) * r
    if axis == "token":
        B, P, E = a.shape
        return torch.stack((a, b), dim=2).reshape(B, 2 * P, E)[:, :n_tokens]
    B, N, P = a.shape
    return torch.stack((a, b), dim=3).reshape(B, N, 2 * P)


def haar1d_mix_ref(x: torch.Tensor, w: torch.Tensor, axis: str) -> torch.Tensor:
    """y = cat(cls, IHaar(w * Haar(x[:, 1:]))) along ``axis`` in {"token", "embed"}.
    w real ``[2, ceil(N/2), E]`` (token) or ``[2, N, E/2]`` (embed), ordered (low, high)."""
    lo, hi = _haar1d_pairs(x[:, 1:], axis)
    y = _ihaar1d_pairs(lo * w[0], hi * w[1], axis, x.shape[1] - 1)
    return torch.cat([x[:, :1], y], dim=1)


def haar1d_mix_bwd_ref(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, axis: str):
    """Closed-form backward: the pair transform is orthogonal and symmetric (and crop^T = pad), so
    dx = IHaar(w * Haar(dy)), dw = sum_b Haar(x_b) * Haar(dy_b)."""
    dlo, dhi = _haar1d_pairs(dy[:, 1:], axis)
    xlo, xhi = _haar1d_pairs(x[:, 1:], axis)
    dxg = _ihaar1d_pairs(dlo * w[0], dhi * w[1], axis, x.shape[1] - 1)
    return torch.cat([dy[:, :1], dxg], dim=1), torch.stack(((xlo * dlo).sum(0), (xhi * dhi).sum(0)))


def haar2d_filterbank(g: torch.Tensor) -> torch.Tensor:
    """The same transform written as the published pywt/pytorch_wavelets Haar
    filter bank (``lo=[1,1]/sqrt2``, ``hi=[1,-1]/sqrt2``, stride 2, no padding
    needed for even sizes under ``mode="zero"``) -- used only to cross-check
    :func:`haar2d`."""
    r = 1.0 / math.sqrt(2.0)
    lo_w = (g[:, :, 0::2] + g[:, :, 1::2]) * r
    hi_w = (g[:, :, 0::2] - g[:, :, 1::2]) * r
    ll = (lo_w[:, 0::2] + lo_w[:, 1::2]) * r
    lh = (lo_w[:, 0::2] - lo_w[:, 1::2]) * r      # low along W, high along H
    hl = (hi_w[:, 0::2] + hi_w[:, 1::2]) * r      # high along W, low along H
    hh = (hi_w[:, 0::2] - hi_w[:, 1::2]) * r
    return torch.stack([ll, lh, hl, hh], dim=1)


# --------------------------------------------------------------------------
# nn.Module occupants of the mix_layer slot, pure torch (CPU baseline / tests)
# --------------------------------------------------------------------------


class OracleFFTMix(nn.Module):
    """Pure-torch occupant of ``mix_layer`` (ctor contract of ``MHPermutMix``,
    ``spectre_vit/models/spectre/layers.py:54``).  Parameter name/shape match
    the product module so state dicts are interchangeable."""

    def __init__(self, embed_dim, token_dim, num_heads, out_channels, *, grid=None):
        super().__init__()
        assert out_channels == embed_dim
        if grid is None:
            s = math.isqrt(token_dim - 1)
            grid = (s, s)
        self.grid = tuple(grid)
        H, W = self.grid
        assert H * W + 1 == token_dim
        f = torch.zeros(H, W // 2 + 1, embed_dim, 2)
        f[..., 0] = 1.0
        self.filter = nn.Parameter(f + 0.02 * torch.randn_like(f))

    def forward(self, x):
        H, W = self.grid
        dt = x.dtype
        return fft_mix_ref(x.float(), self.filter.float(), H, W).to(dt)


class OracleDWTMix(nn.Module):
    def __init__(self, embed_dim, token_dim, num_heads, out_channels, *, grid=None):
        super().__init__()
        assert out_channels == embed_dim
        if grid is None:
            s = math.isqrt(token_dim - 1)
            grid = (s, s)
        self.grid = tuple(grid)
        H, W = self.grid
        assert H * W + 1 == token_dim
        self.weight = nn.Parameter(1.0 + 0.02 * torch.randn(4, (H + 1) // 2, (W + 1) // 2, embed_dim))

    def forward(self, x):
        H, W = self.grid
        dt = x.dtype
        return dwt_mix_ref(x.float(), self.weight.float(), H, W).to(dt)
