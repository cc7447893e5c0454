"""Experiment configs in the reference's flat-module format (``spectre_vit/configs/*.py``) plus one
extra key, ``mixer`` in {"fft", "dwt", "permut"}.  ``NAMED`` maps BASELINE.json's five configurations."""
import os

from .parser import parse_config  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
NAMED = {
    "c1": os.path.join(_HERE, "spectre_vit_cifar100.py"),   # tiny FFT-mix, CIFAR-shaped 32x32/4 -> 8x8
    "c2": os.path.join(_HERE, "spectre_vit_mnist.py"),      # tiny FFT-mix, MNIST-shaped 28x28/4 -> 7x7
    "c3": os.path.join(_HERE, "dwt_cifar100.py"),           # Haar DWT-mix, CIFAR-shaped, bf16, batch 1024
    "c4": os.path.join(_HERE, "vit_s16_fft.py"),            # ViT-S/16 FFT-mix 224^2 -> 14x14, bf16
    "c5": os.path.join(_HERE, "vit_b16_fft_distill.py"),    # ViT-B/16 student <- random-init ViT-B/16 teacher
}


def load_named(name: str):
    return parse_config(NAMED[name])
