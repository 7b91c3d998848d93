#!/bin/bash
# Installs the UNMODIFIED reference package into baseline/_ref (git-ignored; it travels to the GPU
# box with the snapshot) so that bench.py can time the reference's own CPU path there
# (oracle/reference_path.py).  The reference's build backend (poetry-dynamic-versioning) is not in
# this image, so the install runs from a copy under /tmp whose pyproject.toml names setuptools
# instead; the package sources are not touched (diff -r pilott baseline/_ref/pilott: identical).
set -e
REF=${PILOTT_REFERENCE:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
[ -d "$REF/pilott" ] || { echo "no reference at $REF"; exit 0; }
TMP=$(mktemp -d)
cp -r "$REF" "$TMP/ref"
python - "$TMP/ref/pyproject.toml" <<'PY'
import sys
p = sys.argv[1]
s = open(p).read()
s = s[:s.index('[build-system]')] + '''[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "pilott"
version = "0.0.0"

[tool.setuptools.packages.find]
include = ["pilott*"]

[tool.setuptools.package-data]
pilott = ["source/*"]
'''
open(p, 'w').write(s)
PY
rm -rf "$ROOT/baseline/_ref"
python -m pip install -q --no-index --no-build-isolation --no-deps --target "$ROOT/baseline/_ref" "$TMP/ref"
rm -rf "$TMP"
echo "installed $(ls "$ROOT/baseline/_ref")"
