#!/bin/bash
# Installs the UNMODIFIED reference (pure Python) into the git-ignored baseline/_ref, which travels to the GPU box
# with the repo snapshot; tests/test_gpu_e2e_omni.py imports vllm_omni from there. The reference's setup.py imports
# setuptools_scm (absent here, no index): a two-line stand-in that only answers get_version() is put on PYTHONPATH for
# the build; the source tree is read-only, so the build runs on a copy under /tmp. No dependency is installed
# (--no-deps): what vllm_omni's import chain misses at run time is listed in tests/ref_import.py.
set -e
REPO=$(cd "$(dirname "$0")/.." && pwd)
REF=${1:-/root/reference}
[ -d "$REF" ] || { echo "no reference tree at $REF"; exit 0; }
[ -f "$REPO/baseline/_ref/vllm_omni/__init__.py" ] && { echo "baseline/_ref already present"; exit 0; }
TMP=$(mktemp -d)
cp -r "$REF" "$TMP/src"
mkdir -p "$TMP/shim/setuptools_scm"
printf 'def get_version(*a, **k):\n    return "0.0.0+reference.snapshot"\n' > "$TMP/shim/setuptools_scm/__init__.py"
mkdir -p "$REPO/baseline"
(cd "$TMP/src" && VLLM_OMNI_TARGET_DEVICE=cuda PYTHONPATH="$TMP/shim" python -m pip install --no-index --no-build-isolation \
   --no-deps --quiet --target "$REPO/baseline/_ref" .)
rm -rf "$TMP"
echo "installed $(ls "$REPO/baseline/_ref" | tr '\n' ' ')"
