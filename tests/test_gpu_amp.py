"""The slot occupants under the reference's own AMP recipe (`-m gpu`).

``spectre_vit/repl/train.py:205,219,236-238`` trains under ``torch.autocast(dtype=torch.float16)`` with a
``GradScaler``.  ``torch.fft`` is on autocast's fp32 list: inside that region the reference
composition receives an fp32 up-cast of x and returns fp32.  The CUDA mixers must do the same (fp32 kernels, fp32
result) instead of raising, so the training script picks them up with a one-line change."""
import pytest
import torch
import torch.nn as nn

from oracle import mixers_ref as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("mixer", ["fft", "dwt"])
def test_reference_amp_recipe_fp16_autocast_gradscaler(cuda_device, mixer):
    from vit_spectre_experiments_b200 import SpectreViT

    torch.manual_seed(42)
    model = SpectreViT(img_size=32, patch_size=4, in_channels=3, num_classes=10, embed_dim=64, num_encoders=2,
                       num_heads=4, hidden_dim=96, dropout=0.001, activation="gelu", mixer=mixer).to(cuda_device)
    seen = []
    mix0 = model.encoder_blocks.layers[0].mix_layer
    mix0.register_forward_hook(lambda m, inp, out: seen.append((inp[0].detach(), out.detach())))
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.AdamW(model.parameters(), betas=(0.9, 0.999), lr=1e-3, weight_decay=0.01)
    scaler = torch.amp.GradScaler("cuda", enabled=True)                  # train.py:205
    g = torch.Generator().manual_seed(1)
    img = torch.randn(64, 3, 32, 32, generator=g).to(cuda_device)
    label = torch.randint(0, 10, (64,), generator=g).to(cuda_device)
    losses = []
    model.train()
    for _ in range(10):
        with torch.autocast(device_type="cuda", dtype=torch.float16, enabled=True):   # train.py:219
            y_pred = model(img)
        loss = criterion(y_pred, label)
        optimizer.zero_grad(set_to_none=True)                            # train.py:235-238
        scaler.scale(loss).backward()
        scaler.step(optimizer)
        scaler.update()
        losses.append(loss.item())
    assert all(torch.isfinite(torch.tensor(losses))) and losses[-1] < losses[0]
    p = mix0.filter if mixer == "fft" else mix0.weight
    assert p.dtype == torch.float32 and p.grad is not None and p.grad.dtype == torch.float32
    assert torch.isfinite(p.grad).all()

    # the mixer's result inside the autocast region: fp32, and within 1e-5 of the fp32 oracle on the same input
    seen.clear()
    model.eval()
    with torch.no_grad(), torch.autocast(device_type="cuda", dtype=torch.float16):
        model(img)
    x_in, y_out = seen[0]
    assert y_out.dtype == torch.float32
    ref = (O.fft_mix_ref if mixer == "fft" else O.dwt_mix_ref)(x_in.float().cpu(), p.detach().cpu(), 8, 8)
    assert rel_err(y_out, ref) <= 1e-5


def test_half_model_runs_fp32_kernels(cuda_device):
    """``model.half()`` without autocast: fp16 activations are up-cast for the kernel and the result returns as fp16;
    the spectral parameters stay fp32."""
    from vit_spectre_experiments_b200 import SpectralFilterMix

    m = SpectralFilterMix(32, 65, 4, 32).to(cuda_device).half()
    assert m.filter.dtype == torch.float32
    x = torch.randn(3, 65, 32, device=cuda_device).half().requires_grad_(True)
    y = m(x)
    assert y.dtype == torch.float16
    y.float().square().sum().backward()
    assert x.grad.dtype == torch.float16 and m.filter.grad.dtype == torch.float32
    ref = O.fft_mix_ref(x.detach().float().cpu(), m.filter.detach().cpu(), 8, 8)
    assert rel_err(y, ref) <= 2e-3      # fp16 rounding of the result only


def test_permut_backward_after_load_state_dict(cuda_device):
    """the ``permut`` mixer's CUDA backward uses inverse permutations derived from ``perms``; after
    ``load_state_dict`` replaced the permutations in place, dx must follow the NEW ones."""
    from vit_spectre_experiments_b200.model import MHPermutMix

    torch.manual_seed(3)
    a, b = MHPermutMix(16, 10, 2, 16).to(cuda_device), MHPermutMix(16, 10, 2, 16).to(cuda_device)
    x = torch.randn(4, 10, 16, device=cuda_device)
    a(x.clone().requires_grad_(True)).sum().backward()          # builds the cache for a's own permutations
    a.load_state_dict(b.state_dict())
    xa = x.clone().requires_grad_(True)
    ya = a(xa)
    w = torch.randn_like(ya)
    (ya * w).sum().backward()
    xr = x.clone().requires_grad_(True)
    mixed = xr.reshape(4, -1)[:, b.perms] * b.signs              # layers.py:68-70 through torch indexing
    yr = b.linear(mixed.view(4, 10, -1))
    (yr * w).sum().backward()
    assert rel_err(ya, yr) <= 1e-5 and rel_err(xa.grad, xr.grad) <= 1e-5


def test_two_devices_same_process(cuda_device):
    """the opt-in shared-memory attribute is per device; a second GPU driven from the same process must work."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from vit_spectre_experiments_b200 import fft_mix
    from vit_spectre_experiments_b200.fused import norm_act

    outs = []
    for d in ("cuda:0", "cuda:1"):
        g = torch.Generator().manual_seed(0)
        x = torch.randn(4, 197, 384, generator=g).to(d)
        f = torch.randn(14, 8, 384, 2, generator=g).to(d)
        y = fft_mix(x, f, (14, 14))
        y = norm_act(y, torch.ones(384, device=d), torch.zeros(384, device=d), skip=x)
        outs.append(y.cpu())
    assert torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("T,E", [(65, 512), (197, 384), (50, 48)])
def test_fnet_and_fft_module_on_gpu(cuda_device, T, E):
    """``fft_bare`` (orthogonal_permut.py:23-28) and ``modules.FFT`` (modules/spectre.py:5-14) on the GPU: the constant-DFT
    GEMMs against ``torch.fft`` (cuFFT) on the same device, forward and input gradient; fp32 <= 1e-5 of max|ref|
    (full-fp32 GEMMs, TF32 left untouched for the caller), bf16 under autocast <= 3e-2."""
    from vit_spectre_experiments_b200 import FFT, FNetMix

    torch.manual_seed(T + E)
    tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True           # the reference enables it (repl/train.py:36-39)
    try:
        x = torch.randn(6, T, E, device=cuda_device, requires_grad=True)
        w = torch.randn(6, T, E, device=cuda_device)
        m = FNetMix(E, T, 4, E).to(cuda_device)
        y = m(x)
        ref = torch.fft.fft2(x, dim=(-2, -1)).real
        assert rel_err(y, ref) <= 1e-5
        (gy,) = torch.autograd.grad((y * w).sum(), x)
        (gr,) = torch.autograd.grad((ref * w).sum(), x)
        assert rel_err(gy, gr) <= 1e-5
        with torch.autocast("cuda", dtype=torch.bfloat16):
            yb = m(x)
        assert yb.dtype == torch.bfloat16 and rel_err(yb, ref) <= 3e-2
        f = FFT()
        assert rel_err(f(x), torch.fft.rfft(x, dim=-1).real) <= 1e-5
        assert torch.backends.cuda.matmul.allow_tf32 is True
        # the constant matrices are uploaded once per (device, dtype), not per call
        a = m._dft("e", x)
        assert m._dft("e", x) is a
    finally:
        torch.backends.cuda.matmul.allow_tf32 = tf32
