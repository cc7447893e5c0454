"""ctypes binding of ``libspectre_b200.so`` (C ABI: ``include/spectre_b200.h``).

There is deliberately no fallback: if the library has not been built, importing the
kernels raises, and every entry point raises ``RuntimeError`` on a non-zero return code.
"""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, os.environ.get("SB200_LIB_NAME", "libspectre_b200.so"))   # alternate name: A/B builds

# every symbol include/spectre_b200.h declares
EXPORTED_SYMBOLS = (
    "spectre_fftmix_fwd",
    "spectre_fftmix_bwd",
    "spectre_fftmix_bwd_workspace_bytes",
    "spectre_fftmix_bwd_noalloc",
    "spectre_dwtmix_fwd",
    "spectre_dwtmix_bwd",
    "spectre_dwtmix_bwd_workspace_bytes",
    "spectre_dwtmix_ml_fwd",
    "spectre_dwtmix_ml_bwd",
    "spectre_dwtmix_ml_bwd_workspace_bytes",
    "spectre_haar1d_fwd",
    "spectre_haar1d_bwd",
    "spectre_haar1d_bwd_workspace_bytes",
    "spectre_norm_act_fwd",
    "spectre_norm_act_bwd",
    "spectre_norm_act_bwd_workspace_bytes",
    "spectre_norm_act_supported",
    "spectre_patch_gather",
    "spectre_embed_finish_fwd",
    "spectre_embed_finish_bwd",
    "spectre_embed_finish_bwd_workspace_bytes",
    "spectre_permut_fwd",
    "spectre_permut_bwd",
    "spectre_debug_mark_main_kernel_end",
    "spectre_debug_mark_pending",
    "spectre_fftmix_supported",
    "spectre_last_launch_count",
    "spectre_last_error",
    "spectre_version",
)

class NormActDesc(ctypes.Structure):
    """``spectre_norm_act_t`` of include/spectre_b200.h (field order matters)."""

    _fields_ = [
        ("in0", ctypes.c_void_p), ("in1", ctypes.c_void_p), ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
        ("skip", ctypes.c_void_p), ("out", ctypes.c_void_p), ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
        ("dout", ctypes.c_void_p), ("dout2", ctypes.c_void_p), ("d_in0", ctypes.c_void_p), ("d_skip", ctypes.c_void_p),
        ("dgamma", ctypes.c_void_p), ("dbeta", ctypes.c_void_p), ("dcolsum", ctypes.c_void_p), ("workspace", ctypes.c_void_p),
        ("workspace_bytes", ctypes.c_size_t), ("rows", ctypes.c_longlong), ("N", ctypes.c_int), ("Ns", ctypes.c_int),
        ("act_dtype", ctypes.c_int), ("act", ctypes.c_int), ("eps", ctypes.c_float), ("drop_p", ctypes.c_float),
        ("seed", ctypes.c_ulonglong), ("reverse", ctypes.c_int),
        ("out_pool", ctypes.c_void_p), ("d_out_pool", ctypes.c_void_p), ("pool_k", ctypes.c_int),
        ("seed_dev", ctypes.c_void_p),
    ]


_lock = threading.Lock()
_lib = None
_p, _i, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t


def load() -> ctypes.CDLL:
    """Load (once) and type the library.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the spectral mixer has no CPU / torch.fft fallback. "
                "Build it with `python -m vit_spectre_experiments_b200.build` (needs nvcc)."
            )
        lib = ctypes.CDLL(LIB_PATH)
        lib.spectre_dwtmix_ml_fwd.argtypes = [_p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_dwtmix_ml_fwd.restype = _i
        lib.spectre_dwtmix_ml_bwd.argtypes = [_p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_dwtmix_ml_bwd.restype = _i
        lib.spectre_dwtmix_ml_bwd_workspace_bytes.argtypes = [_i, _i, _i, _i]
        lib.spectre_dwtmix_ml_bwd_workspace_bytes.restype = _sz
        lib.spectre_haar1d_fwd.argtypes = [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_haar1d_fwd.restype = _i
        lib.spectre_haar1d_bwd.argtypes = [_p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_haar1d_bwd.restype = _i
        lib.spectre_haar1d_bwd_workspace_bytes.argtypes = [_i, _i, _i, _i]
        lib.spectre_haar1d_bwd_workspace_bytes.restype = _sz
        lib.spectre_debug_mark_main_kernel_end.argtypes = [_p]
        lib.spectre_debug_mark_main_kernel_end.restype = None
        lib.spectre_debug_mark_pending.argtypes = []
        lib.spectre_debug_mark_pending.restype = _i
        lib.spectre_fftmix_fwd.argtypes = [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_fftmix_fwd.restype = _i
        lib.spectre_fftmix_bwd.argtypes = [_p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_fftmix_bwd.restype = _i
        lib.spectre_fftmix_bwd_workspace_bytes.argtypes = [_i, _i, _i, _i]
        lib.spectre_fftmix_bwd_workspace_bytes.restype = _sz
        lib.spectre_fftmix_bwd_noalloc.argtypes = [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_fftmix_bwd_noalloc.restype = _i
        lib.spectre_dwtmix_fwd.argtypes = [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_dwtmix_fwd.restype = _i
        lib.spectre_dwtmix_bwd.argtypes = [_p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_dwtmix_bwd.restype = _i
        lib.spectre_dwtmix_bwd_workspace_bytes.argtypes = [_i, _i, _i, _i]
        lib.spectre_dwtmix_bwd_workspace_bytes.restype = _sz
        lib.spectre_norm_act_fwd.argtypes = [ctypes.POINTER(NormActDesc), _p]
        lib.spectre_norm_act_fwd.restype = _i
        lib.spectre_norm_act_bwd.argtypes = [ctypes.POINTER(NormActDesc), _p]
        lib.spectre_norm_act_bwd.restype = _i
        lib.spectre_norm_act_bwd_workspace_bytes.argtypes = [ctypes.c_longlong, _i]
        lib.spectre_norm_act_bwd_workspace_bytes.restype = _sz
        lib.spectre_norm_act_supported.argtypes = [_i, _i, _i]
        lib.spectre_norm_act_supported.restype = _i
        lib.spectre_patch_gather.argtypes = [_p, _p, _i, _i, _i, _i, _i, _i, _p]
        lib.spectre_patch_gather.restype = _i
        lib.spectre_embed_finish_fwd.argtypes = [_p, _p, _p, _p, _i, _i, _i, _i, ctypes.c_float, ctypes.c_ulonglong, _p, _p]
        lib.spectre_embed_finish_fwd.restype = _i
        lib.spectre_embed_finish_bwd.argtypes = [_p, _p, _p, _p, _sz, _i, _i, _i, _i, ctypes.c_float, ctypes.c_ulonglong, _p, _p]
        lib.spectre_embed_finish_bwd.restype = _i
        lib.spectre_embed_finish_bwd_workspace_bytes.argtypes = [_i, _i, _i]
        lib.spectre_embed_finish_bwd_workspace_bytes.restype = _sz
        lib.spectre_permut_fwd.argtypes = [_p, _p, _p, _p, _i, ctypes.c_longlong, _i, _i, _p]
        lib.spectre_permut_fwd.restype = _i
        lib.spectre_permut_bwd.argtypes = [_p, _p, _p, _p, _i, ctypes.c_longlong, _i, _i, _p]
        lib.spectre_permut_bwd.restype = _i
        lib.spectre_fftmix_supported.argtypes = [_i, _i]
        lib.spectre_fftmix_supported.restype = _i
        lib.spectre_last_launch_count.argtypes = []
        lib.spectre_last_launch_count.restype = _i
        lib.spectre_last_error.argtypes = []
        lib.spectre_last_error.restype = ctypes.c_char_p
        lib.spectre_version.argtypes = []
        lib.spectre_version.restype = ctypes.c_char_p
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().spectre_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


# running count of kernels enqueued through this binding (bench.py reports it)
launch_counter = {"kernels": 0}


def count_launches() -> None:
    launch_counter["kernels"] += load().spectre_last_launch_count()


# Device word the dropout kernels add to their seeds (``seed_dev`` of include/spectre_b200.h).  A training step captured
# in a CUDA graph freezes every host-side seed; train.Trainer(cuda_graph=True) installs an int64 device tensor here and
# bumps it inside the graph, so each replay draws fresh masks (forward and backward of a replay read the same value).
graph_seed = {"tensor": None}


def seed_dev_ptr():
    t = graph_seed["tensor"]
    return None if t is None else t.data_ptr()
