#!/usr/bin/env bash
# Places the UNMODIFIED reference (main.py) under baseline/_ref so that
# `bench.py --impl reference` can import it. The reference ships no setup.py/pyproject,
# so the documented pip command fails on /root/reference itself ("not installable");
# a copy under /tmp gets packaging metadata only (py-modules = ["main"]) and main.py is
# installed byte-for-byte (checked with cmp below). Outcome recorded in DESIGN.md.
set -euo pipefail
cd "$(dirname "$0")/.."
REF=${1:-/root/reference}
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
    --target baseline/_ref "$REF" && exit 0 || echo "direct install failed (expected): packaging a copy"
rm -rf /tmp/ref_pkg && mkdir -p /tmp/ref_pkg
cp "$REF/main.py" "$REF/README.md" /tmp/ref_pkg/
cat > /tmp/ref_pkg/pyproject.toml <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "pytorch-distributed-collective-communication"
version = "0.0.0"
description = "Unmodified reference main.py, packaged only so pip can place it under baseline/_ref"

[tool.setuptools]
py-modules = ["main"]
TOML
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target baseline/_ref /tmp/ref_pkg
cmp baseline/_ref/main.py "$REF/main.py" && echo "baseline/_ref/main.py is identical to the reference"
