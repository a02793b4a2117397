#!/usr/bin/env bash
# Reference-as-run (SURVEY section 8c "Oracle B"): a runnable copy of the reference under the
# git-ignored baseline/_ref/ (it is NOT gpurun-ignored, so it travels to the GPU box).  Never committed; nothing
# of it is product source.
#
#   tools/make_baseline_ref.sh            copy + the seven mechanical edits
#   tools/make_baseline_ref.sh --prebuild ... and compile the reference's four CUDA extensions here for sm_100
#                                         (nvcc cross-compiles without a GPU; ~5 min per translation unit), so the
#                                         GPU box loads them instead of spending ~20 GPU-minutes in the JIT
#
# The only source edit is SURVEY F3: AT_DISPATCH_FLOATING_TYPES(x.type(), ...) does not compile against
# torch 2.11; x.scalar_type() is the documented replacement.  Build flags are the reference's own
# (LBM_propagation_2d.py:11-20: with_cuda=True, extra_cuda_cflags=["--use_fast_math"]).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${JRL_REFERENCE_ROOT:-/root/reference}"
DST="$ROOT/baseline/_ref"
[ -d "$REF/src/LBM" ] || { echo "make_baseline_ref: no reference tree at $REF" >&2; exit 1; }
mkdir -p "$DST/demo"
chmod -R u+w "$DST" 2>/dev/null || true
rm -rf "$DST/src"
cp -r "$REF/src" "$DST/src"
chmod -R u+w "$DST"   # the reference tree is read-only; the copy must not be
cp "$REF"/demo/*.py "$DST/demo/"
cp "$REF/readme.md" "$DST/" 2>/dev/null || true
chmod -R u+w "$DST"
n=$(grep -rl --include=*.cu '\.type()' "$DST/src" | wc -l)
grep -rl --include=*.cu '\.type()' "$DST/src" | xargs sed -i 's/AT_DISPATCH_FLOATING_TYPES(\([a-z_]*\)\.type()/AT_DISPATCH_FLOATING_TYPES(\1.scalar_type()/'
left=$( (grep -rn --include=*.cu 'AT_DISPATCH_FLOATING_TYPES([a-z_]*\.type()' "$DST/src" || true) | wc -l)
echo "make_baseline_ref: copied src/ and demo/*.py to $DST; patched $n .cu files, $left unpatched sites"
[ "$left" -eq 0 ]
if [ "${1:-}" = "--prebuild" ]; then
  python "$ROOT/tools/prebuild_ref_ext.py"
fi
