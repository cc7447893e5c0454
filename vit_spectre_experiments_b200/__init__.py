"""B200-native spectral token mixers behind the Spectre-ViT ``mix_layer`` slot.

Public surface
--------------
``SpectralFilterMix`` / ``HaarDWTMix``   nn.Modules with the ``MHPermutMix`` ctor/forward contract
``FNetMix`` / ``FFT``                    the reference's ``fft_bare`` mixer and ``modules.FFT`` as constant-DFT GEMMs
``Haar1dMix``                            the ``dwt_token`` / ``dwt_embed`` modes the reference docstring names (1-D Haar)
``fft_mix`` / ``dwt_mix``                autograd functions over the C-ABI CUDA library
``SpectreViT`` / ``SpectreEncoderLayer`` host-side mirror of the reference model classes with a
                                         ``mixer=`` switch at the one line that builds ``mix_layer``
"""
from .functional import dwt_mix, dwt_mix_ml, fft_mix, haar1d_mix  # noqa: F401
from .mixers import FFT, FNetMix, Haar1dMix, HaarDWTMix, SpectralFilterMix  # noqa: F401
from .model import (  # noqa: F401
    MHPermutMix,
    SpectralPatchEmbed,
    SpectreEncoder,
    SpectreEncoderLayer,
    SpectreLinear,
    SpectreViT,
)

__version__ = "0.1.0"
