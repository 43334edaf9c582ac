#!/usr/bin/env bash
# Offline install of the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box).
#
# The prescribed command
#   python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
#          --target baseline/_ref /root/reference
# fails: the reference is a script tree with neither setup.py nor pyproject.toml ("Directory
# '/root/reference' is not installable").  As /root/reference is read-only, the python sources are
# copied verbatim to /tmp, a minimal pyproject.toml (packaging metadata only, no source change) is
# added next to them, and that copy is installed with the same pip flags.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/fedmse_ref_XXXX)"
mkdir -p "$TMP/src"
for pkg in Model Trainer Evaluator DataLoader Utils; do
  mkdir -p "$TMP/src/$pkg"
  cp "$REF/src/$pkg"/*.py "$TMP/src/$pkg/"
done
rm -f "$TMP/src/DataLoader/spliter.py"   # a C-style comment block, not python (never imported)
cp "$REF/src/main.py" "$TMP/src/fedmse_reference_main.py"
cat > "$TMP/pyproject.toml" <<'PYP'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "fedmse-decentralized-reference"
version = "0.0.0"
[tool.setuptools]
package-dir = {"" = "src"}
packages = ["Model", "Trainer", "Evaluator", "DataLoader", "Utils"]
py-modules = ["fedmse_reference_main"]
PYP
rm -rf "$REPO/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
       --target "$REPO/baseline/_ref" "$TMP"
rm -rf "$TMP"
echo "installed reference into $REPO/baseline/_ref"
