#!/bin/bash
# One-time offline install of the UNMODIFIED reference into baseline/_ref (git-ignored).
# /root/reference is four stand-alone scripts with no setup.py / pyproject.toml, so
# `pip install /root/reference` answers "not installable" (recorded in DESIGN.md).  Per the
# task's fallback we install from a copy under /tmp to which only a packaging stub is added;
# the four .py files land in baseline/_ref byte-identical to the reference.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
TMP=/tmp/b200dp_ref_copy
rm -rf "$TMP" "$HERE/_ref"
mkdir -p "$TMP"
cp "$SRC"/*.py "$TMP"/
cat > "$TMP/pyproject.toml" <<'EOT'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "torch-distributed-sandbox-reference"
version = "0.0.0"
[tool.setuptools]
py-modules = ["mnist_distributed", "mnist_onegpu", "allreduce_toy", "test_init"]
EOT
python -m pip install --quiet --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$HERE/_ref" "$TMP"
for f in mnist_distributed mnist_onegpu allreduce_toy test_init; do
  cmp "$SRC/$f.py" "$HERE/_ref/$f.py"
done
echo "reference installed unmodified into $HERE/_ref"
