#!/bin/sh
# Recipe for oracle/_ref: the UNMODIFIED reference package (pure Python, BSD-3) installed from where it lies under
# /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot like the built .so files).
# bench.py --impl reference times it (cpu_baseline.kind "reference"); nothing in the product imports it.
# /root/reference is read-only and setup.py writes build/ and egg-info next to itself, so the install runs from a
# temporary copy.  No sources enter the repository history.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${1:-/root/reference}"
[ -d "$REF/seqm" ] || { echo "build_ref: $REF not present (GPU box): keeping the prebuilt oracle/_ref" >&2; exit 0; }
TMP="$(mktemp -d)"
cp -r "$REF" "$TMP/src"
rm -rf "$HERE/_ref"
python -m pip install --quiet --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$HERE/_ref" "$TMP/src" || { echo "build_ref: pip install failed" >&2; rm -rf "$TMP"; exit 1; }
rm -rf "$TMP"
python - "$HERE/_ref" <<'PY'
import os, sys
root = sys.argv[1]
assert os.path.exists(os.path.join(root, 'seqm', 'basics.py')), 'seqm package missing in oracle/_ref'
params = os.path.join(root, 'seqm', 'params')
print('oracle/_ref ready:', root, '| parameter files:', sorted(os.listdir(params)) if os.path.isdir(params) else 'MISSING')
PY
