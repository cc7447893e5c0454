"""Generates ``tests/golden/*.npz`` by IMPORTING THE REFERENCE (read-only at /root/reference) in the
build container.  The GPU box has no /root/reference, so the vectors are committed; re-run with

    python oracle/make_golden.py

Fixtures
  ref_patch_embed.npz   reference ``SpectralPatchEmbed`` (spectre.py:106-156): input, state, output --
                        pins the ``rfft2(norm="ortho")`` convention of the reference's live call site
  ref_fft_module.npz    reference ``FFT`` module (modules/spectre.py:5-14): ``rfft(x, dim=-1).real``
  ref_vit_permut.npz    reference ``SpectreViT`` at HEAD (MHPermutMix in the slot): state + logits --
                        pins the model-class mirror (rows A1, A2, A7, A8)
  ref_vit_fftmix.npz    reference ``SpectreViT`` with ``mix_layer`` := torch.fft composition
                        (rfft2 * filter -> irfft2, the path the kernels replace): state + logits + grads
  ref_vit_dwtmix.npz    same with the Haar DWT occupant
  mixer_fft_<H>x<W>.npz torch.fft composition alone, fwd + autograd bwd, float64
  mixer_dwt_<H>x<W>.npz Haar composition alone, fwd + autograd bwd, float64
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")

from spectre_vit.models.spectre.spectre import SpectralPatchEmbed, SpectreViT  # noqa: E402
from spectre_vit.modules.spectre import FFT  # noqa: E402

from oracle.mixers_ref import OracleDWTMix, OracleFFTMix  # noqa: E402


def npstate(sd):
    return {"sd::" + k: v.detach().cpu().numpy() for k, v in sd.items()}


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(42)  # configs/default.py:1

    emb = SpectralPatchEmbed(embed_dim=16, patch_size=4, num_patches=49, dropout=0.0, in_channels=3).eval()
    with torch.no_grad():
        emb.freq_weight_h.copy_(torch.rand(4) + 0.5)
        emb.freq_weight_w.copy_(torch.rand(3) + 0.5)
    img = torch.randn(2, 3, 28, 28)
    np.savez_compressed(os.path.join(OUT, "ref_patch_embed.npz"), img=img.numpy(), out=emb(img).detach().numpy(),
                        **npstate(emb.state_dict()))

    x = torch.randn(3, 5, 12)
    np.savez_compressed(os.path.join(OUT, "ref_fft_module.npz"), x=x.numpy(), out=FFT()(x).numpy())

    kw = dict(img_size=28, patch_size=4, in_channels=3, num_classes=10, embed_dim=16, num_encoders=2, num_heads=2,
              hidden_dim=24, dropout=0.0)
    vit = SpectreViT(**kw).eval()
    img = torch.randn(2, 3, 28, 28)
    np.savez_compressed(os.path.join(OUT, "ref_vit_permut.npz"), img=img.numpy(), logits=vit(img).detach().numpy(),
                        **npstate(vit.state_dict()))

    for name, cls, kw2 in (("fft", OracleFFTMix, dict(kw)), ("dwt", OracleDWTMix, dict(kw, img_size=32))):
        vit = SpectreViT(**kw2).eval()
        T = (kw2["img_size"] // 4) ** 2 + 1
        for i, layer in enumerate(vit.encoder_blocks.layers):
            torch.manual_seed(100 + i)
            layer.mix_layer = cls(16, T, 2, 16)      # the reference slot, spectre.py:50
            with torch.no_grad():
                for p in layer.mix_layer.parameters():
                    p.add_(0.3 * torch.randn_like(p))
        img = torch.randn(2, 3, kw2["img_size"], kw2["img_size"])
        logits, feat = vit(img, return_features=True)
        loss = logits.square().mean()
        loss.backward()
        grads = {"grad::" + k: p.grad.numpy() for k, p in vit.named_parameters() if "mix_layer" in k}
        np.savez_compressed(os.path.join(OUT, f"ref_vit_{name}mix.npz"), img=img.numpy(),
                            logits=logits.detach().numpy(), feat=feat.detach().numpy(), loss=loss.item(),
                            **npstate(vit.state_dict()), **grads)

    from oracle.mixers_ref import dwt_mix_ref, fft_mix_ref

    for H, W in ((7, 7), (8, 8), (14, 14), (16, 16), (4, 4)):
        g = torch.Generator().manual_seed(H * 100 + W)
        B, E = 3, 6
        x = torch.randn(B, 1 + H * W, E, dtype=torch.float64, generator=g, requires_grad=True)
        f = torch.randn(H, W // 2 + 1, E, 2, dtype=torch.float64, generator=g, requires_grad=True)
        dy = torch.randn(B, 1 + H * W, E, dtype=torch.float64, generator=g)
        y = fft_mix_ref(x, f, H, W)
        y.backward(dy)
        np.savez_compressed(os.path.join(OUT, f"mixer_fft_{H}x{W}.npz"), x=x.detach().numpy(), filt=f.detach().numpy(),
                            dy=dy.numpy(), y=y.detach().numpy(), dx=x.grad.numpy(), dfilt=f.grad.numpy())
    for H, W in ((8, 8), (14, 14), (4, 6)):
        g = torch.Generator().manual_seed(H * 100 + W + 1)
        B, E = 3, 8
        x = torch.randn(B, 1 + H * W, E, dtype=torch.float64, generator=g, requires_grad=True)
        w = torch.randn(4, H // 2, W // 2, E, dtype=torch.float64, generator=g, requires_grad=True)
        dy = torch.randn(B, 1 + H * W, E, dtype=torch.float64, generator=g)
        y = dwt_mix_ref(x, w, H, W)
        y.backward(dy)
        np.savez_compressed(os.path.join(OUT, f"mixer_dwt_{H}x{W}.npz"), x=x.detach().numpy(), w=w.detach().numpy(),
                            dy=dy.numpy(), y=y.detach().numpy(), dx=x.grad.numpy(), dw=w.grad.numpy())
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
