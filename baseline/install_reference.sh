#!/usr/bin/env bash
# Installs the UNMODIFIED reference scripts into baseline/_ref (git-ignored).  The reference ships no setup.py /
# pyproject.toml, so `pip install /root/reference` fails ("not installable"); as the build rules allow, it is
# installed from a copy under /tmp to which only a packaging shim (pyproject.toml listing the three modules) is added.
set -euo pipefail
SRC="${1:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
TMP="$(mktemp -d /tmp/refcopy.XXXX)"
cp "$SRC/client1.py" "$SRC/client2.py" "$SRC/server.py" "$TMP/"
cat > "$TMP/pyproject.toml" <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "ddos-distilbert-fl-reference"
version = "0.0.0"
description = "Packaging shim only: installs the unmodified reference scripts as top-level modules"
[tool.setuptools]
py-modules = ["client1", "client2", "server"]
TOML
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$HERE/_ref" --upgrade "$TMP"
cmp "$HERE/_ref/client1.py" "$SRC/client1.py" && cmp "$HERE/_ref/client2.py" "$SRC/client2.py" && cmp "$HERE/_ref/server.py" "$SRC/server.py" && echo "reference installed, byte-identical"
