#!/usr/bin/env bash
# Test infrastructure (oracle/): build oracle/_ref/libjrl_ref.so from the reference's own
# kernel sources, compiled for the CPU.
#
# The kernel templates are cut out of the .cu files WHERE THEY LIE under the reference
# tree (nothing is copied into this repository: the cut text goes to a scratch directory
# that is removed afterwards, and oracle/_ref/ is git-ignored). "Cut" = keep everything
# above the first host launcher, which in all four files is the first line that starts
# with `void ` (propagation.cu:490, collision_2d.cu:344, macro_compute.cu:159,
# solid_fluid_coupling.cu:254) -- the launchers use torch::Tensor and <<<>>> and are
# replaced by oracle/ref_launch.cpp.
#
# Usage: oracle/build_ref.sh [REFERENCE_ROOT]   (default /root/reference)
# Exits 0 with a message and builds nothing when the reference tree is absent (GPU box).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ref="${1:-/root/reference}"
out="$here/_ref"
src="$ref/src/LBM"
if [ ! -d "$src" ]; then
    echo "build_ref: no reference tree at $ref; keeping any prebuilt $out/libjrl_ref.so"
    exit 0
fi
mkdir -p "$out"
tmp="$(mktemp -d)"
trap 'rm -rf "$tmp"' EXIT
cut_kernels() { sed '/^void /,$d' "$1" > "$2"; }
cut_kernels "$src/LBM_propagation/cuda/propagation.cu"                   "$tmp/ref_prop.inc"
cut_kernels "$src/LBM_collision/cuda/collision_2d.cu"                    "$tmp/ref_collision.inc"
cut_kernels "$src/LBM_macro_compute/cuda/macro_compute.cu"               "$tmp/ref_macro.inc"
cut_kernels "$src/LBM_solid_fluid_coupling/cuda/solid_fluid_coupling.cu" "$tmp/ref_wall.inc"
# -O2 without -ffast-math: IEEE arithmetic. (The reference builds with nvcc --use_fast_math;
# the difference is a few ulp in fp32 divisions and is inside the 1e-5 parity budget.)
g++ -O2 -fopenmp -fPIC -shared -std=c++17 -w \
    -I"$here/shim" -I"$here" -I"$tmp" \
    "$here/ref_launch.cpp" -o "$out/libjrl_ref.so"
echo "build_ref: wrote $out/libjrl_ref.so"
