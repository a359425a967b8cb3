#!/usr/bin/env bash
# Offline install of the UNMODIFIED reference into baseline/_ref (git-ignored).
# The reference's build backend (poetry-core) is not in the image, so the wheel is
# built from a /tmp copy whose pyproject.toml [build-system] points at setuptools;
# no file under p2pfl/ is touched (diff -r /root/reference/p2pfl baseline/_ref/p2pfl
# shows only node.proto, which setuptools does not package, and __pycache__).
set -euo pipefail
cd "$(dirname "$0")/.."
rm -rf /tmp/refcopy baseline/_ref
cp -r /root/reference /tmp/refcopy
python - <<'PY'
p = "/tmp/refcopy/pyproject.toml"
s = open(p).read()
i, j = s.index("[build-system]"), s.index("[tool.ruff]")
s = s[:i] + '[build-system]\nrequires = ["setuptools"]\nbuild-backend = "setuptools.build_meta"\n\n' + s[j:]
s = '[project]\nname = "p2pfl"\nversion = "0.3.0"\n\n[tool.setuptools.packages.find]\ninclude = ["p2pfl*"]\n\n' + s
open(p, "w").write(s)
PY
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref /tmp/refcopy
