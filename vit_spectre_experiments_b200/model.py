"""Host-side mirror of the reference's model classes around the ``mix_layer`` slot.

Same constructor arguments, attribute names and ``state_dict`` keys as the reference, so its
configs / checkpoints / training scripts keep working; the only change is a ``mixer=`` switch at
the single line that builds ``mix_layer``:

    reference                                                     here
    ``SpectreLinear``       models/spectre/layers.py:76-101        :class:`SpectreLinear`
    ``MHPermutMix``         models/spectre/layers.py:53-73         :class:`MHPermutMix`     (mixer="permut")
    ``SpectreEncoderLayer`` models/spectre/spectre.py:29-73        :class:`SpectreEncoderLayer`
    ``SpectreEncoder``      models/spectre/spectre.py:76-103       :class:`SpectreEncoder`
    ``SpectralPatchEmbed``  models/spectre/spectre.py:106-156      :class:`SpectralPatchEmbed`
    ``SpectreViT``          models/spectre/spectre.py:159-202      :class:`SpectreViT`

These are ordinary torch modules (GEMMs, LayerNorm, GELU stay library calls: SURVEY.md section 8f marks
them "next"); the hand-written kernels sit behind ``mix_layer`` only.
"""
from __future__ import annotations

import copy
import itertools
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from .mixers import FNetMix, Haar1dMix, HaarDWTMix, SpectralFilterMix

MIXERS = ("permut", "fft", "dwt", "fnet", "dwt_token", "dwt_embed")

# SB200_FUSED=0 keeps the stock torch composition of LayerNorm / GELU / pool / add around the mixer and
# the GEMMs on CUDA too (A/B measurements); default: the fused norm_act kernels (fused.py).
_FUSED_DEFAULT = os.environ.get("SB200_FUSED", "1") != "0"
_seed_counter = itertools.count(1)


def _dropout_seed() -> int:
    """Per-call seed of the counter-based dropout in the fused tail (deterministic under manual_seed)."""
    return (torch.initial_seed() * 0x9E3779B1 + next(_seed_counter) * 0x85EBCA77) & 0x7FFFFFFFFFFFFFFF


class SpectreLinear(nn.Module):
    """``GELU(LayerNorm(Linear(x))) + skip(x)``; skip is identity when the widths match and an
    adaptive average pool over the channel axis otherwise (layers.py:76-101)."""

    def __init__(self, in_channels: int, out_channels: int, tokens=None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.sparsity = 1
        self.register_buffer("local_idx", torch.arange(0, in_channels, self.sparsity))
        self.local_head = nn.Sequential(
            nn.Linear(len(self.local_idx), out_channels), nn.LayerNorm(out_channels), nn.GELU()
        )
        self.avg_pool = nn.Identity() if in_channels == out_channels else nn.AdaptiveAvgPool1d(out_channels)

    def _kernel_ok(self, x: torch.Tensor) -> bool:
        """CUDA, fused path on, widths the row-wise kernels take, and not inside the reference's fp16 autocast."""
        if not (x.is_cuda and _FUSED_DEFAULT and x.dtype in (torch.float32, torch.bfloat16) and self.sparsity == 1):
            return False
        if torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") != torch.bfloat16:
            return False
        ok = getattr(self, "_na_ok", None)
        if ok is None:
            from . import _lib

            ok = self._na_ok = _lib.load().spectre_norm_act_supported(self.out_channels, self.in_channels, 0) == 1
        return ok

    def forward(self, x: torch.Tensor, dim=-1) -> torch.Tensor:
        if self._kernel_ok(x):
            # Linear (library GEMM) + ONE row-wise kernel for LayerNorm, GELU and the adaptive-pool skip (fp32 stream, as
            # autocast's LayerNorm would give; deterministic backward -- the stock AdaptiveAvgPool1d backward uses atomics)
            from .fused import norm_act

            lin, ln = self.local_head[0], self.local_head[1]
            h = F.linear(x, lin.weight, lin.bias).float()
            return norm_act(h, ln.weight, ln.bias, skip=x.float(), eps=ln.eps, act="gelu")
        return self.local_head(x) + self.avg_pool(x)


class MHPermutMix(nn.Module):
    """The reference's HEAD occupant of the slot (layers.py:53-73): per head a fixed random
    sign flip + permutation of the flattened ``[T*E]`` vector, heads concatenated on the channel
    axis, then ``SpectreLinear(E*heads -> out)``.  Kept so reference configs without a ``mixer``
    key behave as before."""

    def __init__(self, embed_dim: int, token_dim: int, num_heads: int, out_channels: int):
        super().__init__()
        flat = embed_dim * token_dim
        self.num_heads, self.token_dim, self.embed_dim = num_heads, token_dim, embed_dim
        self.concat_dim = embed_dim * num_heads
        signs = torch.randint(0, 2, (num_heads, flat), dtype=torch.float32) * 2 - 1
        self.register_buffer("signs", signs.unsqueeze(0))
        self.register_buffer("perms", torch.stack([torch.randperm(flat) for _ in range(num_heads)]))
        self.linear = SpectreLinear(self.concat_dim, out_channels)

    def _inverse(self) -> torch.Tensor:
        """Inverse permutations for the CUDA backward (plain attribute: not part of ``state_dict``)."""
        # keyed on the buffer's storage, VERSION and device: ``load_state_dict`` copies into ``perms`` in place
        # (same tensor object, bumped version), ``.to(device)`` replaces it
        key = (self.perms.data_ptr(), self.perms._version, self.perms.device)
        if getattr(self, "_inv_key", None) != key:
            from .functional import inverse_permutations

            self._inv, self._inv_key = inverse_permutations(self.perms), key
        return self._inv

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        batch = x.shape[0]
        if x.is_cuda and _FUSED_DEFAULT and x.dtype in (torch.float32, torch.bfloat16) and (self.perms.shape[1] % 4 == 0):
            from .functional import permut_mix   # gather kernel (bit-identical forward, deterministic backward)

            mixed = permut_mix(x, self.perms, self.signs, self._inverse())
        else:
            mixed = x.reshape(batch, -1)[:, self.perms] * self.signs
        return self.linear(mixed.view(batch, self.token_dim, self.concat_dim))


def build_mixer(mixer: str, d_model: int, seq_length: int, nhead: int, grid=None) -> nn.Module:
    """The one line the build changes: what ``SpectreEncoderLayer`` puts in ``mix_layer``
    (reference: ``MHPermutMix(d_model, seq_length, nhead, d_model)``, spectre.py:50)."""
    if callable(mixer):  # any nn.Module factory with the MHPermutMix ctor contract (the slot is open)
        return mixer(d_model, seq_length, nhead, d_model)
    if mixer == "permut":
        return MHPermutMix(d_model, seq_length, nhead, d_model)
    if mixer == "fft":
        return SpectralFilterMix(d_model, seq_length, nhead, d_model, grid=grid)
    if mixer == "dwt":
        return HaarDWTMix(d_model, seq_length, nhead, d_model, grid=grid)
    if mixer == "fnet":
        return FNetMix(d_model, seq_length, nhead, d_model)
    if mixer in ("dwt_token", "dwt_embed"):   # mode names of the reference docstring (spectre.py:33-34)
        return Haar1dMix(d_model, seq_length, nhead, d_model, axis=mixer[4:])
    raise ValueError(f"unknown mixer {mixer!r}; choose from {MIXERS}")


class SpectreEncoderLayer(nn.Module):
    """``x = LN1(mix(x)) + x ; x = LN2(x + FF(x))`` (spectre.py:65-73).

    On CUDA the row-wise pieces around the mixer and the two GEMMs run as fused ``norm_act`` kernels
    (``fused.py``); parameters, ``state_dict`` keys and results are those
    of the torch composition below, which stays the CPU / reference path."""

    def __init__(self, seq_length, d_model, nhead, dim_feedforward, dropout, activation, *, mixer="permut",
                 grid=None, fused=None):
        super().__init__()
        self.mix_layer = build_mixer(mixer, d_model, seq_length, nhead, grid)
        self.linear1 = SpectreLinear(d_model, dim_feedforward)
        self.linear3 = SpectreLinear(dim_feedforward, d_model)
        self.norm1 = nn.LayerNorm(d_model, eps=1e-5)
        self.norm2 = nn.LayerNorm(d_model, eps=1e-5)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.activation = getattr(F, activation) if isinstance(activation, str) else activation
        self.fused = _FUSED_DEFAULT if fused is None else bool(fused)
        self._fusable = None   # resolved on first CUDA use: do the fused row-wise kernels take these widths?

    def _ff_block(self, x: torch.Tensor) -> torch.Tensor:
        return self.dropout2(self.linear3(self.dropout1(self.linear1(x))))

    def _widths_supported(self) -> bool:
        """``spectre_norm_act_supported`` for the three tails of the layer (widths beyond the kernels' limits keep the
        torch composition instead of failing mid-forward)."""
        if self._fusable is None:
            from . import _lib

            lib = _lib.load()
            E, hid = self.norm1.normalized_shape[0], self.linear1.out_channels
            self._fusable = all(lib.spectre_norm_act_supported(n, ns, dt) == 1
                                for dt in (0, 1) for n, ns in ((E, E), (hid, E), (E, hid)))
        return self._fusable

    def use_fused(self, x: torch.Tensor) -> bool:
        if not (self.fused and x.is_cuda and x.dtype in (torch.float32, torch.bfloat16)):
            return False
        if torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") != torch.bfloat16:
            # the reference's fp16 autocast (repl/train.py:219): the stock composition runs around the CUDA mixer,
            # which follows torch.fft's autocast rule (fp32 in, fp32 out)
            return False
        return self._widths_supported()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.use_fused(x):
            return self.forward_fused(x).to(x.dtype)
        x = self.norm1(self.mix_layer(x)) + x
        return self.norm2(x + self._ff_block(x))

    def forward_fused(self, x: torch.Tensor, x_twin=None, return_twin: bool = False):
        """Same arithmetic through the fused kernels.  Under bf16 autocast the residual stream is bf16 from
        end to end (one cast at the encoder entry; every kernel reads and writes bf16, statistics fp32);
        otherwise everything stays float32.  Returns the layer output in the stream dtype.

        With one of the CUDA mixers in the slot the whole layer is ONE autograd node (``fused.encoder_layer``:
        hand-scheduled backward, gradient sums and bias gradients folded into the kernels); any other
        occupant of the slot (``MHPermutMix``, a user module) keeps the per-kernel autograd composition."""
        from .fused import encoder_layer, norm_act

        if torch.is_autocast_enabled() and x.dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        lin1, ln1 = self.linear1.local_head[0], self.linear1.local_head[1]
        lin3, ln3 = self.linear3.local_head[0], self.linear3.local_head[1]
        p1 = self.dropout1.p if self.training else 0.0
        p2 = self.dropout2.p if self.training else 0.0
        s1 = _dropout_seed() if p1 > 0 else 0
        s2 = _dropout_seed() if p2 > 0 else 0
        mix = self.mix_layer
        if isinstance(mix, (SpectralFilterMix, HaarDWTMix)) and x.dim() == 3 and getattr(mix, "levels", 1) == 1:
            kind, param = ("fft", mix.filter) if isinstance(mix, SpectralFilterMix) else ("dwt", mix.weight)
            return encoder_layer(x, kind, param, mix.grid, mix.has_cls, self.norm1, lin1, ln1, lin3, ln3, self.norm2,
                                 p1, p2, s1, s2, x_twin=x_twin, return_twin=return_twin)
        m = mix(x)
        x1 = norm_act(m, self.norm1.weight, self.norm1.bias, skip=x, eps=self.norm1.eps)
        h1 = F.linear(x1, lin1.weight, lin1.bias)
        a1 = norm_act(h1, ln1.weight, ln1.bias, skip=x1, eps=ln1.eps, act="gelu", drop_p=p1, seed=s1)
        h2 = F.linear(a1, lin3.weight, lin3.bias)
        ff = norm_act(h2, ln3.weight, ln3.bias, skip=a1, eps=ln3.eps, act="gelu", drop_p=p2, seed=s2)
        out = norm_act(x1, self.norm2.weight, self.norm2.bias, in1=ff, eps=self.norm2.eps)
        return (out, None) if return_twin else out


class SpectreEncoder(nn.Module):
    """``num_layers`` deep copies of one layer (identical initial parameters, as the reference's
    ``_get_clones``, spectre.py:86) and an outer residual ``output + src`` (spectre.py:103)."""

    def __init__(self, encoder_layer: nn.Module, num_layers: int, norm=None):
        super().__init__()
        self.layers = nn.ModuleList(copy.deepcopy(encoder_layer) for _ in range(num_layers))
        self.num_layers = num_layers
        self.norm = norm

    def forward(self, src: torch.Tensor) -> torch.Tensor:
        out = src
        if all(isinstance(l, SpectreEncoderLayer) and l.use_fused(src) for l in self.layers):
            twin = None                 # second addend of the gradient between stacked one-node layers (fused.py)
            for layer in self.layers:   # bf16 stream under bf16 autocast, fp32 otherwise
                out, twin = layer.forward_fused(out, twin, return_twin=True)
            out = out.to(src.dtype)
        else:
            for layer in self.layers:
                out = layer(out)
        if self.norm is not None:
            out = self.norm(out)
        return out + src


class SpectralPatchEmbed(nn.Module):
    """Non-overlapping PxP patches -> ``Re(rfft2(patch, norm="ortho"))`` -> learnable per-frequency
    gains -> Linear -> prepend CLS, add positional embeddings, dropout (spectre.py:106-156)."""

    def __init__(self, embed_dim, patch_size, num_patches, dropout, in_channels):
        super().__init__()
        self.P = patch_size
        self.embed_dim = embed_dim
        half = patch_size // 2 + 1
        self.freq_weight_h = nn.Parameter(torch.ones(patch_size))
        self.freq_weight_w = nn.Parameter(torch.ones(half))
        self.proj = nn.Linear(in_channels * patch_size * half, embed_dim)
        self.cls_token = nn.Parameter(torch.randn(1, 1, embed_dim))
        self.position_embeddings = nn.Parameter(torch.randn(1, num_patches + 1, embed_dim))
        self.dropout = nn.Dropout(dropout)

    def forward(self, img: torch.Tensor) -> torch.Tensor:
        if img.is_cuda and _FUSED_DEFAULT:
            return self.forward_folded(img)
        batch, chans = img.shape[:2]
        P = self.P
        patches = img.unfold(2, P, P).unfold(3, P, P)                       # [B, C, nH, nW, P, P]
        n = patches.size(2) * patches.size(3)
        patches = patches.contiguous().view(batch, chans, n, P, P)
        spec = torch.fft.rfft2(patches.float(), norm="ortho").real          # [B, C, N, P, P//2+1]
        spec = spec * self.freq_weight_h.view(1, 1, 1, P, 1) * self.freq_weight_w.view(1, 1, 1, 1, -1)
        tokens = self.proj(spec.permute(0, 2, 1, 3, 4).flatten(2))          # [B, N, E]
        tokens = torch.cat((self.cls_token.expand(batch, -1, -1), tokens), dim=1)
        return self.dropout(tokens + self.position_embeddings)

    def _cos_dft(self, device) -> torch.Tensor:
        """``D[(h, w), (kh, kw)] = cos(2 pi (kh h + kw w) / P) / P``: ``Re(rfft2(p, norm="ortho")) = p_flat @ D``."""
        D = getattr(self, "_D", None)
        if D is None or D.device != device:
            P, half = self.P, self.P // 2 + 1
            h = torch.arange(P, dtype=torch.float64)
            ang = (h.view(P, 1, 1, 1) * h.view(1, 1, P, 1) + h.view(1, P, 1, 1) * h[:half].view(1, 1, 1, half))
            D = (torch.cos(2 * torch.pi * ang / P) / P).reshape(P * P, P * half).float().to(device)
            self._D = D   # plain attribute: not a buffer, so state_dict keys stay the reference's
        return D

    def _folded_matrix(self, device, chans: int) -> torch.Tensor:
        """``M [C*P*P, E]`` from the three parameters.  Under autograd it is rebuilt every call (the graph must reach the
        parameters); without (evaluation, ``torch.no_grad``) it is cached until a parameter changes (in-place updates
        bump ``_version``; ``load_state_dict`` copies in place)."""
        P, half = self.P, self.P // 2 + 1
        params = (self.proj.weight, self.freq_weight_h, self.freq_weight_w)
        cacheable = not (torch.is_grad_enabled() and any(p.requires_grad for p in params))
        key = (device, chans) + tuple((p.data_ptr(), p._version) for p in params)
        if cacheable and getattr(self, "_M_key", None) == key:
            return self._M
        with torch.autocast("cuda", enabled=False):
            wf = (self.proj.weight.float().view(self.embed_dim, chans, P, half)
                  * self.freq_weight_h.float().view(1, 1, P, 1) * self.freq_weight_w.float().view(1, 1, 1, half))
            # [C, P*P, P*half] @ [C, P*half, E] -> [C*P*P, E]
            M = torch.matmul(self._cos_dft(device), wf.view(self.embed_dim, chans, P * half).permute(1, 2, 0))
            M = M.reshape(chans * P * P, self.embed_dim)
        if cacheable:
            self._M, self._M_key = M.detach(), key
        return M

    def forward_folded(self, img: torch.Tensor) -> torch.Tensor:
        """The same map as ONE GEMM: ``Re(rfft2)``, the frequency gains and the
        projection are all linear in the patch, so ``tokens = patch_flat @ M + b`` with
        ``M[(c,h,w), e] = sum_{kh,kw} D[(h,w),(kh,kw)] fwh[kh] fww[kw] Wp[e,(c,kh,kw)]`` rebuilt from the
        parameters every call (42 MFLOP at P=16; autograd reaches all three parameters through it).  Replaces the
        cuFFT launch, its complex -> real copy, two broadcast multiplies and a permute copy over a
        ``[B, C, N, P, P/2+1]`` fp32 tensor by one patch-gather copy + one GEMM."""
        batch, chans = img.shape[:2]
        P, half = self.P, self.P // 2 + 1
        nh, nw = img.shape[2] // P, img.shape[3] // P
        M = self._folded_matrix(img.device, chans)
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled() else img.dtype
        if (dt in (torch.float32, torch.bfloat16) and P % 4 == 0 and img.shape[3] % 4 == 0 and self.embed_dim % 8 == 0):
            # patch gather + cast | GEMM | cls / position / dropout pass: the CUDA kernels of csrc/embed.cu
            from .functional import patch_embed

            p = self.dropout.p if self.training else 0.0
            return patch_embed(img, M, self.proj.bias, self.cls_token, self.position_embeddings, P, p,
                               _dropout_seed() if p > 0 else 0, dt)
        patches = img[:, :, :nh * P, :nw * P].reshape(batch, chans, nh, P, nw, P).permute(0, 2, 4, 1, 3, 5)
        flat = patches.to(dt).reshape(batch * nh * nw, chans * P * P)        # one gather copy (+ cast)
        tokens = torch.addmm(self.proj.bias.to(dt), flat, M.to(dt)).view(batch, nh * nw, self.embed_dim)
        tokens = torch.cat((self.cls_token.expand(batch, -1, -1).to(tokens.dtype), tokens), dim=1)
        return self.dropout(tokens + self.position_embeddings)


class SpectreViT(nn.Module):
    """embed -> encoder -> ``mlp_head(x[:, 0])``; ``return_features=True`` also returns the CLS
    token (spectre.py:159-202).  ``mixer`` selects the ``mix_layer`` occupant."""

    def __init__(self, img_size=32, patch_size=4, in_channels=3, num_classes=10, embed_dim=768, num_encoders=12,
                 num_heads=12, hidden_dim=3072, dropout=0.1, activation="gelu", *, mixer="permut"):
        super().__init__()
        side = img_size // patch_size
        num_patches = side * side
        self.mixer = mixer
        self.embeddings_block = SpectralPatchEmbed(embed_dim, patch_size, num_patches, dropout, in_channels)
        layer = SpectreEncoderLayer(
            seq_length=num_patches + 1, d_model=embed_dim, nhead=num_heads, dim_feedforward=hidden_dim,
            dropout=dropout, activation=activation, mixer=mixer, grid=(side, side),
        )
        self.encoder_blocks = SpectreEncoder(layer, num_layers=num_encoders)
        self.mlp_head = nn.Sequential(SpectreLinear(embed_dim, num_classes))

    def forward(self, x: torch.Tensor, return_features: bool = False):
        x = self.encoder_blocks(self.embeddings_block(x))
        cls_token = x[:, 0, :]
        logits = self.mlp_head(cls_token)
        return (logits, cls_token) if return_features else logits
