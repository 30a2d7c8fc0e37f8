#!/bin/bash
# Offline install of the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box) for
# `bench.py --impl reference` and its cpu_baseline.measured leg.  /root/reference is read-only, so the build runs
# from a copy under /tmp; the wheelhouse has no setuptools_scm, so a 3-line stand-in that reports a fixed version
# sits on PYTHONPATH during the build (nothing of it is installed).
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=/tmp/tmx_ref_src; SHIM=/tmp/tmx_ref_shim
rm -rf "$SRC" "$SHIM" "$ROOT/baseline/_ref"
cp -r /root/reference "$SRC"
mkdir -p "$SHIM/setuptools_scm"
printf 'def get_version(*a, **k):\n    return "0.0.0+trianglemix"\n' > "$SHIM/setuptools_scm/__init__.py"
cd "$SRC"
PYTHONPATH="$SHIM" VLLM_TARGET_DEVICE=empty python -m pip install --no-index --no-build-isolation --no-deps \
  --find-links /opt/wheelhouse --target "$ROOT/baseline/_ref" "$SRC" 2>&1 | tail -3
for f in vllm/attention/trianglemix_attention.py vllm/attention/ascend_trianglemix.py vllm/attention/trianglemix_config.py; do
  cmp "/root/reference/$f" "$ROOT/baseline/_ref/$f" && echo "identical: $f"
done
