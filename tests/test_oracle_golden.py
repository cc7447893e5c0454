"""CPU tests: the oracle against the golden vectors generated from the reference itself
(oracle/make_golden.py) and against algebraic properties of the path."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import mixers_ref as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return np.load(os.path.join(GOLD, name))


def test_golden_files_present():
    names = {os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.npz"))}
    for need in ("ref_patch_embed.npz", "ref_fft_module.npz", "ref_vit_permut.npz", "ref_vit_fftmix.npz",
                 "ref_vit_dwtmix.npz", "mixer_fft_7x7.npz", "mixer_fft_8x8.npz", "mixer_fft_14x14.npz",
                 "mixer_dwt_8x8.npz"):
        assert need in names


def test_direct_dft_matches_reference_patch_embed():
    """The oracle's library-free float64 DFT reproduces the output of the reference's own
    SpectralPatchEmbed (spectre.py:124-156) => the rfft2/ortho convention is pinned."""
    g = load("ref_patch_embed.npz")
    img = g["img"].astype(np.float64)
    B, C = img.shape[:2]
    P = 4
    n = img.shape[2] // P
    patches = img.reshape(B, C, n, P, n, P).transpose(0, 1, 2, 4, 3, 5).reshape(B, C, n * n, P, P)
    spec = O.rfft2_ortho_np(patches).real
    spec = spec * g["sd::freq_weight_h"].reshape(1, 1, 1, P, 1) * g["sd::freq_weight_w"].reshape(1, 1, 1, 1, -1)
    tok = spec.transpose(0, 2, 1, 3, 4).reshape(B, n * n, -1) @ g["sd::proj.weight"].T.astype(np.float64) + g["sd::proj.bias"]
    out = np.concatenate([np.broadcast_to(g["sd::cls_token"], (B, 1, tok.shape[-1])), tok], 1) + g["sd::position_embeddings"]
    assert np.abs(out - g["out"]).max() < 5e-6


def test_direct_dft_matches_reference_fft_module():
    g = load("ref_fft_module.npz")
    x = g["x"].astype(np.float64)
    n = x.shape[-1]
    F = np.exp(-2j * np.pi * np.outer(np.arange(n), np.arange(n // 2 + 1)) / n)
    assert np.abs((x @ F).real - g["out"]).max() < 2e-6


@pytest.mark.parametrize("hw", ["4x4", "7x7", "8x8", "14x14", "16x16"])
def test_fft_mixer_golden(hw):
    """torch.fft composition (golden, fp64) == float64 direct-DFT restatement == closed-form backward."""
    g = load(f"mixer_fft_{hw}.npz")
    H, W = map(int, hw.split("x"))
    x, f, dy = (torch.from_numpy(g[k]) for k in ("x", "filt", "dy"))
    fc = torch.view_as_complex(f).numpy()
    assert np.abs(O.fft_mix_np(g["x"], fc, H, W) - g["y"]).max() < 1e-12
    assert (O.fft_mix_ref(x, f, H, W) - torch.from_numpy(g["y"])).abs().max() < 1e-12
    dx, df = O.fft_mix_bwd_ref(dy, x, f, H, W)
    assert (dx - torch.from_numpy(g["dx"])).abs().max() < 1e-12
    assert (df - torch.from_numpy(g["dfilt"])).abs().max() < 1e-11


@pytest.mark.parametrize("hw", ["8x8", "14x14", "4x6"])
def test_dwt_mixer_golden(hw):
    g = load(f"mixer_dwt_{hw}.npz")
    H, W = map(int, hw.split("x"))
    x, w, dy = (torch.from_numpy(g[k]) for k in ("x", "w", "dy"))
    assert (O.dwt_mix_ref(x, w, H, W) - torch.from_numpy(g["y"])).abs().max() < 1e-13
    dx, dw = O.dwt_mix_bwd_ref(dy, x, w, H, W)
    assert (dx - torch.from_numpy(g["dx"])).abs().max() < 1e-12
    assert (dw - torch.from_numpy(g["dw"])).abs().max() < 1e-12


def test_irfft2_ignores_imag_of_dc_and_nyquist_columns():
    """torch.fft.irfft2 semantics for a NON-Hermitian product."""
    rng = np.random.default_rng(0)
    for H, W in ((7, 7), (8, 8), (14, 14)):
        Z = rng.standard_normal((2, H, W // 2 + 1)) + 1j * rng.standard_normal((2, H, W // 2 + 1))
        ref = torch.fft.irfft2(torch.from_numpy(Z), s=(H, W), norm="ortho").numpy()
        assert np.abs(O.irfft2_ortho_np(Z, H, W) - ref).max() < 1e-13


def test_identity_parseval_linearity_haar_pr():
    torch.manual_seed(0)
    for H, W in ((7, 7), (8, 8), (14, 14)):
        x = torch.randn(2, 1 + H * W, 4, dtype=torch.float64)
        ident = torch.zeros(H, W // 2 + 1, 4, 2, dtype=torch.float64)
        ident[..., 0] = 1
        assert (O.fft_mix_ref(x, ident, H, W) - x).abs().max() < 1e-13
        X = O.rfft2_ortho_np(x[:, 1:].reshape(2, H, W, 4).permute(0, 3, 1, 2).numpy())
        c = np.full(W // 2 + 1, 2.0)
        c[0] = 1
        if W % 2 == 0:
            c[-1] = 1
        assert abs((np.abs(X) ** 2 * c).sum() - (x[:, 1:] ** 2).sum().item()) < 1e-9
        f = torch.randn(H, W // 2 + 1, 4, 2, dtype=torch.float64)
        assert (O.fft_mix_ref(2 * x, f, H, W)[:, 1:] - 2 * O.fft_mix_ref(x, f, H, W)[:, 1:]).abs().max() < 1e-12
    g = torch.randn(2, 8, 6, 3, dtype=torch.float64)
    assert (O.ihaar2d(O.haar2d(g)) - g).abs().max() < 1e-14
    assert (O.haar2d(g) - O.haar2d_filterbank(g)).abs().max() < 1e-14
    assert abs(O.haar2d(g).pow(2).sum() - g.pow(2).sum()) < 1e-10


def test_closed_form_backward_equals_autograd():
    torch.manual_seed(1)
    for H, W in ((7, 7), (8, 8), (14, 14), (6, 5)):
        x = torch.randn(2, 1 + H * W, 4, dtype=torch.float64, requires_grad=True)
        f = torch.randn(H, W // 2 + 1, 4, 2, dtype=torch.float64, requires_grad=True)
        y = O.fft_mix_ref(x, f, H, W)
        dy = torch.randn_like(y)
        y.backward(dy)
        dx, df = O.fft_mix_bwd_ref(dy, x.detach(), f.detach(), H, W)
        assert (dx - x.grad).abs().max() < 1e-12 and (df - f.grad).abs().max() < 1e-12


@pytest.mark.parametrize("H,W", [(7, 7), (5, 8), (13, 3)])
def test_oracle_dwt_odd_grid_zero_mode(H, W):
    """Odd grids under ``mode="zero"`` (dwt_experiments.py:56): pad one zero row / column, transform, crop.
    Unit weights reconstruct the input; the closed-form backward equals autograd of the forward."""
    import torch

    from oracle import mixers_ref as O

    torch.manual_seed(4)
    B, E = 3, 8
    x = torch.randn(B, 1 + H * W, E, dtype=torch.float64, requires_grad=True)
    h2, w2 = (H + 1) // 2, (W + 1) // 2
    ones = torch.ones(4, h2, w2, E, dtype=torch.float64)
    assert (O.dwt_mix_ref(x, ones, H, W) - x).abs().max() < 1e-12
    w = (1 + 0.3 * torch.randn(4, h2, w2, E, dtype=torch.float64)).requires_grad_(True)
    dy = torch.randn(B, 1 + H * W, E, dtype=torch.float64)
    y = O.dwt_mix_ref(x, w, H, W)
    gx, gw = torch.autograd.grad(y, (x, w), dy)
    dx, dw = O.dwt_mix_bwd_ref(dy, x.detach(), w.detach(), H, W)
    assert (dx - gx).abs().max() < 1e-12 and (dw - gw).abs().max() < 1e-12


def test_oracle_multilevel_haar_layout_and_recursion():
    """The in-place multi-level Haar of the oracle: (i) one level == the single-level oracle under the documented
    de-interleaving; (ii) J levels == recursing the single-level transform on the LL band (the Mallat pyramid the
    reference's DWTForward(J=3) computes); (iii) inverse undoes it; (iv) closed-form backward == autograd."""
    import torch

    from oracle import mixers_ref as O

    torch.manual_seed(8)
    B, H, W, E = 2, 8, 8, 4
    g = torch.randn(B, H, W, E, dtype=torch.float64)
    one = O.haar_ml_inplace(g, 1)
    s4 = O.haar2d(g)                                   # [B, 4, H/2, W/2, E]: LL, lowW/highH, highW/lowH, HH
    assert torch.allclose(one[:, 0::2, 0::2], s4[:, 0]) and torch.allclose(one[:, 1::2, 0::2], s4[:, 1])
    assert torch.allclose(one[:, 0::2, 1::2], s4[:, 2]) and torch.allclose(one[:, 1::2, 1::2], s4[:, 3])
    three = O.haar_ml_inplace(g, 3)
    ll1 = s4[:, 0]
    s4b = O.haar2d(ll1)                                # level 2 on the LL band
    assert torch.allclose(three[:, 2::4, 0::4], s4b[:, 1]) and torch.allclose(three[:, 0::4, 2::4], s4b[:, 2])
    assert torch.allclose(three[:, 2::4, 2::4], s4b[:, 3])
    s4c = O.haar2d(s4b[:, 0])                          # level 3
    assert torch.allclose(three[:, 0::8, 0::8], s4c[:, 0]) and torch.allclose(three[:, 4::8, 4::8], s4c[:, 3])
    assert torch.allclose(three[:, 1::2, 1::2], s4[:, 3])          # level-1 details untouched by later levels
    assert torch.allclose(O.haar_ml_inplace(three, 3, inverse=True), g)
    assert abs(three.pow(2).sum() - g.pow(2).sum()) < 1e-9         # orthonormal
    x = torch.randn(B, 1 + H * W, E, dtype=torch.float64, requires_grad=True)
    w = (1 + 0.3 * torch.randn(H, W, E, dtype=torch.float64)).requires_grad_(True)
    dy = torch.randn(B, 1 + H * W, E, dtype=torch.float64)
    gx, gw = torch.autograd.grad(O.dwt_ml_mix_ref(x, w, H, W, 3), (x, w), dy)
    dx, dw = O.dwt_ml_mix_bwd_ref(dy, x.detach(), w.detach(), H, W, 3)
    assert (dx - gx).abs().max() < 1e-12 and (dw - gw).abs().max() < 1e-12
