"""CPU tests of the host side: codelet arithmetic (host build of the device header), the C-ABI
library's exported symbols, the model-class mirror against reference-generated golden vectors,
module contract (state_dict, deepcopy, dtype policy), loud failure without a GPU."""
import copy
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
PKG = os.path.join(ROOT, "vit_spectre_experiments_b200")


@pytest.fixture(scope="module")
def codelets(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("codelets") / "libcodelets.so")
    subprocess.run(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-I", os.path.join(PKG, "csrc"), "-o", so,
                    os.path.join(ROOT, "tests", "host_codelets.cpp")], check=True)
    return ctypes.CDLL(so)


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 16])
def test_codelets_match_numpy(codelets, n):
    rng = np.random.default_rng(n)
    z = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    for d in (-1, 1):
        re, im = z.real.copy(), z.imag.copy()
        assert codelets.codelet_cfft(n, d, _p(re), _p(im)) == 0
        ref = np.fft.fft(z.astype(np.complex128)) if d < 0 else np.fft.ifft(z.astype(np.complex128)) * n
        assert np.abs(re + 1j * im - ref).max() / np.abs(ref).max() < 5e-7
    x = rng.standard_normal(n).astype(np.float32)
    re, im = np.zeros(n // 2 + 1, np.float32), np.zeros(n // 2 + 1, np.float32)
    codelets.codelet_rfft(n, _p(x), _p(re), _p(im))
    ref = np.fft.rfft(x.astype(np.float64))
    assert np.abs(re + 1j * im - ref).max() / np.abs(ref).max() < 5e-7
    # C2R on a NON-Hermitian half spectrum: imag of DC / Nyquist must be ignored like torch.fft.irfft
    Z = rng.standard_normal(n // 2 + 1) + 1j * rng.standard_normal(n // 2 + 1)
    re, im = Z.real.astype(np.float32), Z.imag.astype(np.float32)
    out = np.zeros(n, np.float32)
    codelets.codelet_irfft(n, _p(re), _p(im), _p(out))
    ref = torch.fft.irfft(torch.tensor(re.astype(np.float64) + 1j * im.astype(np.float64)), n=n, norm="forward").numpy()
    assert np.abs(out - ref).max() / np.abs(ref).max() < 5e-7


def test_library_exports_every_declared_symbol():
    from vit_spectre_experiments_b200 import _lib

    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "spectre_b200.h")).read()
    declared = set(re.findall(r"\b(spectre_[a-z_0-9]+)\s*\(", header))
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert b"sm_100a" in lib.spectre_version()
    assert lib.spectre_fftmix_supported(14, 14) == 1 and lib.spectre_fftmix_supported(7, 7) == 1
    assert lib.spectre_fftmix_supported(8, 8) == 1 and lib.spectre_fftmix_supported(5, 9) == 0
    # argument validation happens before any CUDA call, so it is testable without a GPU
    assert lib.spectre_fftmix_fwd(None, None, None, 1, 8, 8, 8, 1, 0, None) == -1
    assert b"null" in lib.spectre_last_error()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.addressof(buf)
    p += (-p) % 16
    assert lib.spectre_fftmix_fwd(p, p, p, 1, 8, 8, 7, 1, 0, None) == -2      # odd E
    assert lib.spectre_fftmix_fwd(p, p, p, 1, 8, 8, 8, 1, 5, None) == -4      # dtype
    assert lib.spectre_fftmix_fwd(p + 4, p, p, 1, 8, 8, 8, 1, 0, None) == -5  # alignment
    assert lib.spectre_dwtmix_fwd(p, p, p, 1, 7, 7, 7, 1, 0, None) == -2      # odd E (odd grids are zero-padded)


def test_model_mirror_matches_reference_golden():
    """SpectreViT mirror + reference state_dict (HEAD MHPermutMix occupant) => reference logits."""
    from vit_spectre_experiments_b200 import SpectreViT

    g = np.load(os.path.join(GOLD, "ref_vit_permut.npz"))
    m = SpectreViT(img_size=28, patch_size=4, in_channels=3, num_classes=10, embed_dim=16, num_encoders=2,
                   num_heads=2, hidden_dim=24, dropout=0.0, mixer="permut").eval()
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd::")}
    assert set(sd) == set(m.state_dict())
    m.load_state_dict(sd, strict=True)
    with torch.no_grad():
        out = m(torch.from_numpy(g["img"]))
    assert (out - torch.from_numpy(g["logits"])).abs().max() < 1e-6


def test_patch_embed_mirror_matches_reference_golden():
    from vit_spectre_experiments_b200 import SpectralPatchEmbed

    g = np.load(os.path.join(GOLD, "ref_patch_embed.npz"))
    emb = SpectralPatchEmbed(16, 4, 49, 0.0, 3).eval()
    emb.load_state_dict({k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd::")}, strict=True)
    with torch.no_grad():
        assert (emb(torch.from_numpy(g["img"])) - torch.from_numpy(g["out"])).abs().max() < 1e-6


def test_mixer_module_contract():
    from vit_spectre_experiments_b200 import HaarDWTMix, SpectralFilterMix, SpectreViT

    m = SpectralFilterMix(48, 50, 8, 48)
    assert m.grid == (7, 7) and tuple(m.filter.shape) == (7, 4, 48, 2) and m.filter.dtype == torch.float32
    assert list(m.state_dict()) == ["filter"]
    m2 = copy.deepcopy(m)
    assert torch.equal(m2.filter, m.filter) and m2.filter is not m.filter
    m.bfloat16()
    assert m.filter.dtype == torch.float32          # learned spectral parameters stay fp32
    d = HaarDWTMix(512, 65, 16, 512)
    assert tuple(d.weight.shape) == (4, 4, 4, 512) and list(d.state_dict()) == ["weight"]
    assert tuple(HaarDWTMix(48, 50, 8, 48).weight.shape) == (4, 4, 4, 48)   # 7x7 grid: zero-padded to 8x8
    with pytest.raises(ValueError):
        SpectralFilterMix(48, 51, 8, 48)            # 50 tokens is not a square grid + CLS
    vit = SpectreViT(img_size=32, patch_size=4, num_classes=100, embed_dim=64, num_encoders=3, num_heads=4,
                     hidden_dim=96, dropout=0.0, mixer="fft")
    keys = list(vit.state_dict())
    assert "encoder_blocks.layers.2.mix_layer.filter" in keys
    # clones start identical, like the reference's _get_clones (spectre.py:86)
    assert torch.equal(vit.encoder_blocks.layers[0].mix_layer.filter, vit.encoder_blocks.layers[2].mix_layer.filter)


def test_no_cpu_fallback():
    from vit_spectre_experiments_b200 import SpectralFilterMix

    m = SpectralFilterMix(8, 65, 2, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.randn(2, 65, 8))


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                # nor the installed reference (baseline/_ref): only bench.py builds the reference model / teacher from it
                assert "baseline" not in src and "spectre_vit." not in src.replace("spectre_vit/", ""), f


@pytest.mark.parametrize("T,E", [(65, 64), (50, 48), (197, 24)])
def test_fnet_mix_matches_fft2_real(T, E):
    """``fft_bare`` occupant of the slot (orthogonal_permut.py:23-28): two real GEMMs == Re(fft2) over (T, E),
    forward and input gradient (float64 so the comparison is at rounding level)."""
    from vit_spectre_experiments_b200.mixers import FNetMix

    torch.manual_seed(3)
    m = FNetMix(E, T, 4, E)
    assert len(m.state_dict()) == 0                       # constant matrices stay out of checkpoints
    x = torch.randn(3, T, E, dtype=torch.float64, requires_grad=True)
    y = m(x)
    ref = torch.fft.fft2(x, dim=(-2, -1)).real
    assert y.shape == x.shape
    assert (y - ref).abs().max() <= 1e-9 * ref.abs().max()
    g = torch.randn_like(y)
    (dx,) = torch.autograd.grad(y, x, g)
    (dref,) = torch.autograd.grad(ref, x, g)
    assert (dx - dref).abs().max() <= 1e-9 * dref.abs().max()


def test_fnet_mixer_in_the_model():
    from vit_spectre_experiments_b200.model import SpectreViT

    torch.manual_seed(0)
    net = SpectreViT(img_size=32, patch_size=4, in_channels=3, num_classes=10, embed_dim=32, num_encoders=2,
                     num_heads=4, hidden_dim=48, dropout=0.0, activation="gelu", mixer="fnet")
    out = net(torch.randn(2, 3, 32, 32))
    assert out.shape == (2, 10) and torch.isfinite(out).all()
    out.sum().backward()


def test_permut_inverse_cache_follows_load_state_dict():
    """``load_state_dict`` copies into ``perms`` in place; the cached inverse permutations (CUDA backward of
    the ``permut`` mixer) must be rebuilt, not reused."""
    from vit_spectre_experiments_b200.functional import inverse_permutations
    from vit_spectre_experiments_b200.model import MHPermutMix

    torch.manual_seed(0)
    a, b = MHPermutMix(8, 5, 2, 8), MHPermutMix(8, 5, 2, 8)
    assert not torch.equal(a.perms, b.perms)
    assert torch.equal(a._inverse(), inverse_permutations(a.perms))
    a.load_state_dict(b.state_dict())
    assert torch.equal(a.perms, b.perms)
    assert torch.equal(a._inverse(), inverse_permutations(b.perms))
    first = a._inverse()
    assert a._inverse() is first            # unchanged buffer: the cache is used


def test_mixer_construction_validates_grid_and_widths():
    """unsupported grids fail at construction with a clear message; layer widths the fused row-wise kernels
    cannot take are detected by the ``spectre_norm_act_supported`` query (the layer then keeps the torch composition)."""
    import pytest

    from vit_spectre_experiments_b200 import HaarDWTMix, SpectralFilterMix, _lib

    with pytest.raises(ValueError, match="no CUDA kernel"):
        SpectralFilterMix(16, 26, 2, 16)                      # 5x5 grid: not compiled
    SpectralFilterMix(16, 50, 2, 16)                          # 7x7 is
    with pytest.raises(ValueError, match="H\\*W <= 256"):
        HaarDWTMix(16, 1 + 32 * 32, 2, 16, levels=2)
    lib = _lib.load()
    assert lib.spectre_norm_act_supported(384, 1536, 1) == 1 and lib.spectre_norm_act_supported(1536, 384, 0) == 1
    assert lib.spectre_norm_act_supported(5120, 5120, 0) == 0 and lib.spectre_norm_act_supported(12, 12, 0) == 0
    assert lib.spectre_norm_act_supported(512, 768, 1) == 1 and lib.spectre_norm_act_supported(48, 256, 0) == 1


def test_fft_module_matches_reference_golden():
    """``modules.FFT`` (spectre_vit/modules/spectre.py:5-14: ``rfft(x, dim=-1).real``) as a constant cosine-DFT GEMM
    reproduces the output the REFERENCE's own module produced (tests/golden/ref_fft_module.npz) and torch.fft."""
    import numpy as np

    from vit_spectre_experiments_b200 import FFT

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_fft_module.npz"))
    m = FFT()
    assert len(m.state_dict()) == 0
    out = m(torch.from_numpy(g["x"]))
    assert out.shape == g["out"].shape and np.abs(out.numpy() - g["out"]).max() < 2e-6
    x = torch.randn(4, 9, 50, dtype=torch.float64, requires_grad=True)
    y = m(x)
    ref = torch.fft.rfft(x, dim=-1).real
    assert (y - ref).abs().max().item() < 1e-12
    w = torch.randn_like(ref)
    (gy,) = torch.autograd.grad((y * w).sum(), x)
    (gr,) = torch.autograd.grad((ref * w).sum(), x)
    assert (gy - gr).abs().max().item() < 1e-12


def test_bench_reference_arm_contract():
    """``bench.py --impl reference`` (the CPU arm the driver runs beside the GPU arm): one JSON line with the contract's
    keys; ``kind`` is "reference" when the unmodified reference is installed under baseline/_ref, else "port"."""
    import json
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--config", "c2",
                        "--cpu-batch", "16", "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["impl"] == "reference" and d["unit"] == "images/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    have_ref = os.path.isdir(os.path.join(root, "baseline", "_ref", "spectre_vit"))
    assert cb["kind"] == ("reference" if have_ref else "port") and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert d["config"]["workload"].startswith("Spectre-ViT tiny FFT-mix, MNIST")


def test_reference_class_with_oracle_mixer_matches_mirror():
    """With baseline/_ref present: the reference's OWN SpectreViT (oracle mixer in its slot, as bench.py's reference arm
    builds it) and this repo's mirror with the same state produce the same logits on CPU."""
    import pytest

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.isdir(os.path.join(root, "baseline", "_ref", "spectre_vit")):
        pytest.skip("baseline/_ref not installed (run baseline/install_ref.sh)")
    import sys
    from types import SimpleNamespace

    sys.path.insert(0, root)
    import bench
    from oracle.mixers_ref import OracleFFTMix
    from vit_spectre_experiments_b200 import SpectreViT

    cfg = SimpleNamespace(img_size=28, patch_size=4, in_channels=3, num_classes=10, embed_dim=16, num_encoders=2,
                          num_heads=2, hidden_dim=24, dropout=0.0, activation="gelu", mixer="fft")
    torch.manual_seed(0)
    ref = bench.reference_model(cfg).eval()
    mir = SpectreViT(img_size=28, patch_size=4, in_channels=3, num_classes=10, embed_dim=16, num_encoders=2, num_heads=2,
                     hidden_dim=24, dropout=0.0, mixer=lambda d, t, h, o: OracleFFTMix(d, t, h, o)).eval()
    mir.load_state_dict(ref.state_dict(), strict=True)
    x = torch.randn(3, 3, 28, 28)
    with torch.no_grad():
        a, b = ref(x), mir(x)
    assert (a - b).abs().max().item() <= 1e-5 * a.abs().max().item()
