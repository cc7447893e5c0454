#!/bin/bash
# Installs the UNMODIFIED reference (pure Python: spectre_vit + its dinov3 submodule's model code) into baseline/_ref
# (git-ignored, travels to the GPU box) so that
#   * bench.py --impl reference times the reference's OWN SpectreViT class (oracle mixer in its mix_layer slot), and
#   * config C5's teacher is the reference's dinov3_vitb16(pretrained=False) + DinoClassifier.
# `pip install /root/reference` fails as is (flat layout with two top-level packages and requires-python >= 3.13,
# pyproject.toml:6), so the sources are copied to /tmp and given packaging metadata only; no source file is edited.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${1:-/root/reference}
TMP=$(mktemp -d /tmp/refcopy.XXXXXX)
mkdir -p "$TMP/pkg"
cp -r "$REF/spectre_vit" "$TMP/pkg/spectre_vit"
cp -r "$REF/dinov3/dinov3" "$TMP/pkg/dinov3"
find "$TMP/pkg" -name "__pycache__" -prune -exec rm -rf {} +
cat > "$TMP/pkg/setup.py" <<'PY'
from setuptools import find_namespace_packages, setup

setup(name="spectralvit-reference", version="0.1.0", python_requires=">=3.10",
      packages=find_namespace_packages(include=["spectre_vit*", "dinov3*"]),
      package_data={"": ["*.yaml", "*.yml", "*.json", "*.txt"]})
PY
rm -rf "$HERE/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP/pkg" 2>&1 | tail -3
rm -rf "$TMP"
python - <<PY
import sys
sys.path.insert(0, "$HERE/_ref")
from spectre_vit.models.spectre.spectre import SpectreViT          # noqa: F401
from spectre_vit.distillation import DinoClassifier                 # noqa: F401
from dinov3.hub.backbones import dinov3_vitb16                      # noqa: F401
print("baseline/_ref: reference importable")
PY
