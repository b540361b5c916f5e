#!/usr/bin/env bash
# Installs the unmodified reference into baseline/_ref (git-ignored) so that `bench.py --impl reference` and the
# tests/test_reference_*.py parity tests can import it.  The reference tree has neither setup.py nor pyproject.toml,
# so it is copied to a scratch directory and given a packaging shim there; the installed .py files are byte-identical.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
SRC="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/nidt_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
cp -r "$SRC"/. "$TMP"/
chmod -R u+w "$TMP"
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup, find_namespace_packages
setup(name="neuroimagedisttraining-ref", version="0.0",
      packages=find_namespace_packages(include=["fedml_api*", "fedml_core*", "fedml_experiments*"]))
PY
rm -rf "$ROOT/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$ROOT/baseline/_ref" "$TMP"
echo "installed $(find "$ROOT/baseline/_ref" -name '*.py' | wc -l) reference sources under baseline/_ref"
