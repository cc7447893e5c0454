"""1-D Haar mixers ("dwt_token" / "dwt_embed", mode names of spectre_vit/models/spectre/spectre.py:33-34; the reference
has no code for them, so the operator is this repo's -- parity unpinned).  CPU: the oracle's closed-form backward against
autograd, reconstruction and energy.  `-m gpu`: the CUDA kernels (through the C ABI) against the oracle."""
import pytest
import torch

from oracle import mixers_ref as O


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def wshape(N, E, axis):
    return (2, (N + 1) // 2, E) if axis == "token" else (2, N, E // 2)


@pytest.mark.parametrize("axis", ["token", "embed"])
@pytest.mark.parametrize("N", [64, 49, 1])
def test_oracle_backward_is_autograd_of_forward(axis, N):
    g = torch.Generator().manual_seed(3 + N)
    B, E = 3, 16
    x = torch.randn(B, 1 + N, E, generator=g, dtype=torch.float64, requires_grad=True)
    w = torch.randn(*wshape(N, E, axis), generator=g, dtype=torch.float64, requires_grad=True)
    dy = torch.randn(B, 1 + N, E, generator=g, dtype=torch.float64)
    O.haar1d_mix_ref(x, w, axis).backward(dy)
    dx, dw = O.haar1d_mix_bwd_ref(dy, x.detach(), w.detach(), axis)
    assert rel_err(dx, x.grad) < 1e-12 and rel_err(dw, w.grad) < 1e-12


@pytest.mark.parametrize("axis", ["token", "embed"])
def test_oracle_unit_weights_reconstruct_and_preserve_energy(axis):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 1 + 50, 24, generator=g, dtype=torch.float64)
    w = torch.ones(*wshape(50, 24, axis), dtype=torch.float64)
    assert rel_err(O.haar1d_mix_ref(x, w, axis), x) < 1e-13
    lo, hi = O._haar1d_pairs(x[:, 1:], axis)
    assert abs((lo.pow(2).sum() + hi.pow(2).sum()) / x[:, 1:].pow(2).sum() - 1) < 1e-12


def test_build_mixer_knows_the_reference_mode_names():
    from vit_spectre_experiments_b200.model import MIXERS, build_mixer

    assert "dwt_token" in MIXERS and "dwt_embed" in MIXERS
    m = build_mixer("dwt_token", 48, 50, 4)
    assert tuple(m.weight.shape) == (2, 25, 48) and m.axis == "token"   # 49 tokens without CLS -> 25 pairs
    m = build_mixer("dwt_embed", 48, 50, 4)
    assert tuple(m.weight.shape) == (2, 49, 24) and m.axis == "embed"
    with pytest.raises(ValueError):
        build_mixer("dwt_embed", 44, 50, 4)


CASES = [
    # B, N, E, dtype
    (8, 64, 512, torch.float32),        # C1 / C3 token counts (8x8 grid)
    (67, 49, 48, torch.float32),        # C2: odd token count (zero token appended), E = 48
    (33, 196, 384, torch.bfloat16),     # C4 ViT-S
    (5, 196, 768, torch.bfloat16),      # C5 ViT-B
    (300, 7, 8, torch.bfloat16),        # more images than slabs, one channel vector
    (2, 1, 16, torch.float32),          # a single token
]


@pytest.mark.gpu
@pytest.mark.parametrize("B,N,E,dtype", CASES)
@pytest.mark.parametrize("axis", ["token", "embed"])
@pytest.mark.parametrize("has_cls", [True, False])
def test_haar1d_fwd_bwd(cuda_device, B, N, E, dtype, axis, has_cls):
    from vit_spectre_experiments_b200 import haar1d_mix

    g = torch.Generator().manual_seed(17 + B + N + E)
    T = N + int(has_cls)
    x = torch.randn(B, T, E, generator=g).to(dtype)
    dy = torch.randn(B, T, E, generator=g).to(dtype)
    w = 1.0 + 0.5 * torch.randn(*wshape(N, E, axis), generator=g)
    xd, wd = x.to(cuda_device).requires_grad_(True), w.to(cuda_device).requires_grad_(True)
    y = haar1d_mix(xd, wd, has_cls, axis)
    assert y.dtype == dtype and y.shape == x.shape
    y.backward(dy.to(cuda_device))
    xr, dyr = x.float(), dy.float()
    if not has_cls:      # the oracle carries a CLS row: prepend a dummy one
        xr = torch.cat([torch.zeros(B, 1, E), xr], 1)
        dyr = torch.cat([torch.zeros(B, 1, E), dyr], 1)
    y_ref = O.haar1d_mix_ref(xr, w, axis)
    dx_ref, dw_ref = O.haar1d_mix_bwd_ref(dyr, xr, w, axis)
    if not has_cls:
        y_ref, dx_ref = y_ref[:, 1:], dx_ref[:, 1:]
    tol = 1e-5 if dtype == torch.float32 else 3e-2
    assert rel_err(y, y_ref.to(dtype)) <= tol
    assert rel_err(xd.grad, dx_ref.to(dtype)) <= tol
    assert wd.grad.dtype == torch.float32 and rel_err(wd.grad, dw_ref) <= (2e-5 if dtype == torch.float32 else 2e-3)
    if has_cls:
        assert torch.equal(y[:, 0].cpu(), x[:, 0]) and torch.equal(xd.grad[:, 0].cpu(), dy[:, 0])


@pytest.mark.gpu
@pytest.mark.parametrize("mixer", ["dwt_token", "dwt_embed"])
def test_haar1d_in_the_encoder_slot(cuda_device, mixer):
    """The mode names build a layer that trains: unit weights are the identity mixer, fp16 autocast follows the fp32 rule."""
    from vit_spectre_experiments_b200 import SpectreEncoderLayer

    torch.manual_seed(0)
    layer = SpectreEncoderLayer(65, 64, 4, 128, 0.0, "gelu", mixer=mixer).to(cuda_device)
    x = torch.randn(6, 65, 64, device=cuda_device)
    with torch.no_grad():
        layer.mix_layer.weight.fill_(1.0)
        assert rel_err(layer.mix_layer(x), x) <= 1e-6
    out = layer(x)
    out.square().mean().backward()
    assert torch.isfinite(out).all() and torch.isfinite(layer.mix_layer.weight.grad).all()
    with torch.autocast("cuda", dtype=torch.float16):
        assert layer.mix_layer(x.half()).dtype == torch.float32
    with torch.autocast("cuda", dtype=torch.bfloat16):
        assert layer.mix_layer(x).dtype == torch.bfloat16
