"""``parse_config`` with the reference's semantics (``spectre_vit/configs/parser.py:9-27``): a config is a
flat Python module; public names become attributes; a ``__base__ = "file.py"`` key pulls in a sibling
module whose values are merged ON TOP of the child's (the reference's in-place ``mod |= base_mod``, i.e.
the base wins on a clash -- kept as is).  Unlike the reference it also runs on Python < 3.13
(``SimpleNamespace(mapping)`` positional form) and accepts a filesystem path or a dotted module name.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
from types import SimpleNamespace


def _module_to_dict(module) -> dict:
    return {k: getattr(module, k) for k in dir(module) if not k.startswith("_") or k == "__base__"}


def _import(path_or_name: str):
    if path_or_name.endswith(".py") and os.path.exists(path_or_name):
        spec = importlib.util.spec_from_file_location("_spectre_cfg_" + str(abs(hash(path_or_name))), path_or_name)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    return importlib.import_module(path_or_name.replace("/", ".").removesuffix(".py"))


def parse_config(config_path: str) -> SimpleNamespace:
    cfg = _module_to_dict(_import(config_path))
    base = cfg.get("__base__")
    if base:
        if config_path.endswith(".py") and os.path.exists(config_path):
            base_ref = os.path.join(os.path.dirname(config_path), base)
        else:
            dotted = config_path.replace("/", ".").removesuffix(".py")
            base_ref = ".".join(dotted.split(".")[:-1] + [base.removesuffix(".py")])
        cfg |= _module_to_dict(_import(base_ref))
    return SimpleNamespace(**cfg)
