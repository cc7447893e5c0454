"""``nn.Module`` occupants of the reference's ``mix_layer`` slot backed by the CUDA kernels.

Contract (kept from ``MHPermutMix``, ``spectre_vit/models/spectre/layers.py:53-73``): constructed
positionally as ``(embed_dim, token_dim, num_heads, out_channels)`` by
``SpectreEncoderLayer.__init__`` (``spectre.py:50``), called as ``forward(x:[B, token_dim, embed_dim])
-> [B, token_dim, out_channels]`` (``spectre.py:66``), deep-copied by ``_get_clones``
(``spectre.py:86``), moved with ``.to()``, and serialised through ``state_dict()`` -- so the module
holds only Parameters and plain ints.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .functional import dwt_mix, dwt_mix_ml, fft_mix, haar1d_mix, haar1d_weight_shape


def _resolve_grid(token_dim: int, grid, has_cls: bool):
    n = token_dim - int(has_cls)
    if grid is None:
        s = math.isqrt(n)
        grid = (s, s)
    H, W = int(grid[0]), int(grid[1])
    if H * W != n:
        raise ValueError(f"grid {H}x{W} does not cover token_dim={token_dim} (has_cls={has_cls})")
    return H, W


def _require_kernel(kind: str, H: int, W: int, E: int) -> None:
    """Fail at construction, not at the first forward, when no kernel is compiled for the grid (there is no
    fallback to torch.fft).  Skipped when the library has not been built yet: the forward raises then anyway."""
    if E % 2:
        raise ValueError(f"{kind} mixer: embed_dim must be even, got {E}")
    try:
        from . import _lib

        lib = _lib.load()
    except (RuntimeError, OSError):
        return
    if kind == "fft" and lib.spectre_fftmix_supported(H, W) != 1:
        raise ValueError(f"SpectralFilterMix: no CUDA kernel is compiled for a {H}x{W} grid "
                         "(see fftmix_supported in csrc/fftmix.cu for the list)")


class _Fp32ParamMixin:
    """Learned spectral parameters stay float32 under ``.half()`` / ``.bfloat16()`` / ``.to(dtype)``:
    the kernels take fp32 parameters and emit fp32 parameter gradients whatever the activation dtype."""

    _fp32_param = ""

    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        p = getattr(self, self._fp32_param)
        if p.dtype != torch.float32:
            p.data = p.data.float()
            if p.grad is not None:
                p.grad.data = p.grad.data.float()
        return self

    @staticmethod
    def _activation(x: torch.Tensor):
        """Dtype the kernels run in, following what ``torch.fft`` does in the same place.

        * bf16 autocast (the build's training stream): bf16 kernels, bf16 result;
        * any other autocast dtype -- the reference trains under ``torch.autocast(dtype=torch.float16)`` + ``GradScaler``
          (``spectre_vit/repl/train.py:205,219``): ``torch.fft`` is on autocast's fp32 list there, so ``x`` is up-cast
          to float32, the fp32 kernels run and the result stays float32;
        * no autocast: fp32 / bf16 tensors run as they are; a float16 / float64 tensor (``model.half()``) runs the fp32
          kernels and the result is cast back to its dtype.
        Returns ``(x_for_the_kernel, dtype_of_the_result)``."""
        if torch.is_autocast_enabled(x.device.type):
            dt = torch.get_autocast_dtype(x.device.type)
            if dt == torch.bfloat16:
                return (x if x.dtype == dt else x.to(dt)), dt
            return x.float(), torch.float32
        if x.dtype in (torch.float32, torch.bfloat16):
            return x, x.dtype
        return x.float(), x.dtype


class SpectralFilterMix(_Fp32ParamMixin, nn.Module):
    """rfft2 over the patch grid -> learned complex filter -> irfft2, CLS passed through.

    Parameter ``filter``: float32 ``[H, W//2+1, E, 2]`` (viewed complex).  ``num_heads`` is accepted
    for signature compatibility and unused.
    """

    _fp32_param = "filter"

    def __init__(self, embed_dim: int, token_dim: int, num_heads: int, out_channels: int, *, grid=None,
                 has_cls: bool = True, norm: str = "ortho", init_std: float = 0.02):
        super().__init__()
        if out_channels != embed_dim:
            raise ValueError("SpectralFilterMix keeps the channel count: out_channels must equal embed_dim")
        if norm != "ortho":
            raise ValueError('only norm="ortho" (the reference convention, spectre.py:136) is implemented')
        self.embed_dim, self.token_dim, self.num_heads = embed_dim, token_dim, num_heads
        self.has_cls = bool(has_cls)
        self.grid = _resolve_grid(token_dim, grid, self.has_cls)
        H, W = self.grid
        _require_kernel("fft", H, W, embed_dim)
        f = torch.zeros(H, W // 2 + 1, embed_dim, 2)
        f[..., 0] = 1.0
        self.filter = nn.Parameter(f + init_std * torch.randn_like(f))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        xa, out_dt = self._activation(x)
        y = fft_mix(xa, self.filter, self.grid, self.has_cls)
        return y if y.dtype == out_dt else y.to(out_dt)

    def extra_repr(self) -> str:
        return f"embed_dim={self.embed_dim}, grid={self.grid}, has_cls={self.has_cls}"


class HaarDWTMix(_Fp32ParamMixin, nn.Module):
    """Single-level 2-D Haar DWT over the patch grid -> learned real sub-band weights -> IDWT.

    Parameter ``weight``: float32 ``[4, ceil(H/2), ceil(W/2), E]`` ordered (LL, W-low/H-high, W-high/H-low, HH).
    Odd grids follow ``mode="zero"`` (dwt_experiments.py:56): zero row / column appended, result cropped.
    ``levels > 1``: multi-level transform, ``weight`` float32 ``[H, W, E]`` with the coefficients addressed in place on
    the token lattice (include/spectre_b200.h, ``spectre_dwtmix_ml_fwd``).
    """

    _fp32_param = "weight"

    def __init__(self, embed_dim: int, token_dim: int, num_heads: int, out_channels: int, *, grid=None,
                 has_cls: bool = True, levels: int = 1, init_std: float = 0.02):
        super().__init__()
        if out_channels != embed_dim:
            raise ValueError("HaarDWTMix keeps the channel count: out_channels must equal embed_dim")
        self.embed_dim, self.token_dim, self.num_heads = embed_dim, token_dim, num_heads
        self.has_cls = bool(has_cls)
        self.levels = int(levels)
        self.grid = _resolve_grid(token_dim, grid, self.has_cls)
        H, W = self.grid
        if self.levels == 1:
            self.weight = nn.Parameter(1.0 + init_std * torch.randn(4, (H + 1) // 2, (W + 1) // 2, embed_dim))
        else:
            # J levels (the reference call site uses J=3): coefficients in place on the token lattice, [H, W, E]
            if self.levels < 1 or H % (1 << self.levels) or W % (1 << self.levels):
                raise ValueError(f"levels={levels} needs H and W to be multiples of 2^levels, got {H}x{W}")
            if H * W > 256 or embed_dim % 4:
                raise ValueError(f"multi-level Haar kernel: needs H*W <= 256 and embed_dim % 4 == 0, got {H}x{W}, E={embed_dim}")
            self.weight = nn.Parameter(1.0 + init_std * torch.randn(H, W, embed_dim))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        xa, out_dt = self._activation(x)
        if self.levels == 1:
            y = dwt_mix(xa, self.weight, self.grid, self.has_cls)
        else:
            y = dwt_mix_ml(xa, self.weight, self.grid, self.has_cls, self.levels)
        return y if y.dtype == out_dt else y.to(out_dt)

    def extra_repr(self) -> str:
        return f"embed_dim={self.embed_dim}, grid={self.grid}, has_cls={self.has_cls}"


class Haar1dMix(_Fp32ParamMixin, nn.Module):
    """Single-level 1-D Haar DWT along the token axis (``axis="token"``: the reference's ``dwt_token`` mode name) or the
    embedding axis (``"embed"``: ``dwt_embed``) -> learned real weights per coefficient -> inverse (spectre.py:33-34 name
    the modes; the reference has no code for them, the operator follows the 2-D ``HaarDWTMix``).

    Parameter ``weight``: float32 ``[2, ceil(N/2), E]`` (token axis; N tokens without CLS, an odd N is zero-padded) or
    ``[2, N, E/2]`` (embedding axis), ordered (low, high).  ``MHPermutMix`` ctor / forward contract."""

    _fp32_param = "weight"

    def __init__(self, embed_dim: int, token_dim: int, num_heads: int, out_channels: int, *, axis: str = "token",
                 has_cls: bool = True, init_std: float = 0.02):
        super().__init__()
        if out_channels != embed_dim:
            raise ValueError("Haar1dMix keeps the channel count: out_channels must equal embed_dim")
        if axis not in ("token", "embed"):
            raise ValueError(f"axis must be 'token' or 'embed', got {axis!r}")
        if embed_dim % 8:
            raise ValueError(f"Haar1dMix kernels need embed_dim % 8 == 0, got {embed_dim}")
        self.embed_dim, self.token_dim, self.num_heads = embed_dim, token_dim, num_heads
        self.axis, self.has_cls = axis, bool(has_cls)
        n = token_dim - int(self.has_cls)
        if n < 1:
            raise ValueError(f"token_dim={token_dim} leaves no tokens to mix")
        self.weight = nn.Parameter(1.0 + init_std * torch.randn(*haar1d_weight_shape(n, embed_dim, axis)))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        xa, out_dt = self._activation(x)
        y = haar1d_mix(xa, self.weight, self.has_cls, self.axis)
        return y if y.dtype == out_dt else y.to(out_dt)

    def extra_repr(self) -> str:
        return f"embed_dim={self.embed_dim}, token_dim={self.token_dim}, axis={self.axis}, has_cls={self.has_cls}"


def _cos_sin(n: int):
    """float64 ``cos`` / ``sin`` of ``2 pi j k / n`` with the angle reduced exactly (integer ``j k mod n``)."""
    k = torch.arange(n, dtype=torch.int64)
    ang = (2.0 * math.pi / n) * ((k[:, None] * k[None, :]) % n).to(torch.float64)
    return torch.cos(ang), torch.sin(ang)


class _DftBuffers:
    """Constant DFT matrices kept as float64 masters on the CPU and cached per (device, dtype): the cast / upload happens
    once, not on every forward, and nothing is added to ``state_dict()``."""

    def _dft(self, name: str, like: torch.Tensor) -> torch.Tensor:
        cache = self.__dict__.setdefault("_dft_cache", {})
        key = (name, like.device, like.dtype)
        m = cache.get(key)
        if m is None:
            m = cache[key] = self._dft_master[name].to(device=like.device, dtype=like.dtype)
        return m

    def __deepcopy__(self, memo):      # _get_clones deep-copies the layer: share the masters, drop the device cache
        import copy

        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k == "_dft_cache":
                continue
            new.__dict__[k] = v if k == "_dft_master" else copy.deepcopy(v, memo)
        return new


class _NoTF32:
    """fp32 GEMMs in full fp32 inside the block (the reference's ``torch.fft`` is fp32-exact; TF32 would cost 1e-3),
    whatever the caller's ``allow_tf32`` (the reference enables it, repl/train.py:36-39); restored on exit."""

    def __enter__(self):
        self.prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False

    def __exit__(self, *exc):
        torch.backends.cuda.matmul.allow_tf32 = self.prev
        return False


class _ConstMatmul(torch.autograd.Function):
    """``x @ c`` (``left=False``) or ``c @ x`` (``left=True``) against a constant matrix ``c``, forward AND backward in
    full fp32 for fp32 operands (autograd would run the backward GEMM later, under the caller's TF32 setting)."""

    @staticmethod
    def forward(ctx, x, c, left):
        ctx.save_for_backward(c)
        ctx.left = left
        with _NoTF32():
            return torch.matmul(c, x) if left else torch.matmul(x, c)

    @staticmethod
    def backward(ctx, g):
        (c,) = ctx.saved_tensors
        with _NoTF32():
            gx = torch.matmul(c.transpose(-1, -2), g) if ctx.left else torch.matmul(g, c.transpose(-1, -2))
        return gx, None, None


def _matmul_exact(a: torch.Tensor, b: torch.Tensor, const_left: bool = False) -> torch.Tensor:
    """``a @ b`` where one operand is a constant DFT matrix (``b``, or ``a`` with ``const_left``)."""
    if const_left:
        return _ConstMatmul.apply(b, a, True)
    return _ConstMatmul.apply(a, b, False)


class FFT(_DftBuffers, nn.Module):
    """``torch.fft.rfft(x, dim=-1).real`` (``spectre_vit/modules/spectre.py:5-14``): ``[..., E] -> [..., E//2 + 1]``.

    The real part of the real DFT is one dense cosine matrix, ``x @ C`` with ``C[e, k] = cos(2 pi e k / E)``: GEMM-shaped,
    so it is a library GEMM against a constant (no cuFFT plan, no complex intermediate, differentiable through
    ``matmul``); fp32 inputs multiply in full fp32.  Parameter-free, as the reference module."""

    def __init__(self) -> None:
        super().__init__()
        self._dft_master = {}

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        E = x.shape[-1]
        name = f"rfft_cos_{E}"
        if name not in self._dft_master:
            self._dft_master[name] = _cos_sin(E)[0][:, : E // 2 + 1].contiguous()
        dt = x.dtype if x.dtype in (torch.float32, torch.float64, torch.bfloat16, torch.float16) else torch.float32
        return _matmul_exact(x.to(dt), self._dft(name, x.to(dt)))


class FNetMix(_DftBuffers, nn.Module):
    """``fft_bare`` (FNet): ``Re(fft2(x, dim=(-2, -1)))`` over (tokens, channels), parameter-free
    (``spectre_vit/repl/orthogonal_permut.py:23-28``; mode name ``spectre.py:31``).

    ``Re(F_T x F_E) = C_T x C_E - S_T x S_E`` with the cosine / sine DFT matrices, i.e. two dense real GEMMs
    ``x @ [C_E | S_E]`` and ``[C_T | -S_T] @ [xC ; xS]``: GEMM-shaped work, left to the library GEMM (cuBLAS) --
    no custom kernel, no complex intermediates, and it runs in the activation dtype under autocast (fp32 inputs in
    full fp32).  The matrices are float64 masters cached per (device, dtype): nothing is added to ``state_dict()``
    and no conversion runs per forward."""

    def __init__(self, embed_dim: int, token_dim: int, num_heads: int, out_channels: int):
        super().__init__()
        if out_channels != embed_dim:
            raise ValueError("FNetMix keeps the channel count: out_channels must equal embed_dim")
        self.embed_dim, self.token_dim, self.num_heads = embed_dim, token_dim, num_heads
        ce, se = _cos_sin(embed_dim)
        ct, st = _cos_sin(token_dim)
        self._dft_master = {"e": torch.cat([ce, se], dim=1), "t": torch.cat([ct, -st], dim=1)}   # [E, 2E], [T, 2T]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 3 or x.shape[1] != self.token_dim or x.shape[2] != self.embed_dim:
            raise ValueError(f"FNetMix expects [B, {self.token_dim}, {self.embed_dim}], got {tuple(x.shape)}")
        E = self.embed_dim
        if torch.is_autocast_enabled(x.device.type):
            x = x.to(torch.get_autocast_dtype(x.device.type))
        xe = _matmul_exact(x, self._dft("e", x))                       # [B, T, 2E] = [x C_E | x S_E]
        stacked = torch.cat([xe[..., :E], xe[..., E:]], dim=1)         # [B, 2T, E]
        return _matmul_exact(self._dft("t", x), stacked, const_left=True)   # C_T (x C_E) - S_T (x S_E)

    def extra_repr(self) -> str:
        return f"embed_dim={self.embed_dim}, token_dim={self.token_dim}"
