#!/usr/bin/env bash
# Installs the UNMODIFIED reference (zhc180/distributed-llm-trainer, mounted at /root/reference) into baseline/_ref
# for `bench.py --impl reference`.  The reference ships no setup.py / pyproject.toml, so it is installed from a copy
# with a packaging shim that maps src/{models,data,training,eval,utils} to top-level packages -- exactly what its own
# scripts do with sys.path.insert.  Source files stay byte-identical (checked with diff -r below).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/refcopy.XXXXXX)"
cp -r "$REF"/. "$TMP"/
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup
pk = ["models", "data", "training", "eval", "utils"]
import os
pk = [p for p in pk if os.path.isdir(os.path.join("src", p))]
for p in pk:
    init = os.path.join("src", p, "__init__.py")
    if not os.path.exists(init):
        open(init, "w").close()
setup(name="distributed-llm-trainer-reference", version="0", packages=pk, package_dir={p: f"src/{p}" for p in pk})
PY
rm -rf "$ROOT/baseline/_ref"
mkdir -p "$ROOT/baseline"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --no-deps --target "$ROOT/baseline/_ref" "$TMP" -q
for p in models data training eval utils; do
  [ -d "$REF/src/$p" ] && diff -r -x '__pycache__' -x '__init__.py' "$REF/src/$p" "$ROOT/baseline/_ref/$p" && echo "identical: $p"
done
rm -rf "$TMP"
