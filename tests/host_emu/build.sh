#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY: compile the kernel sources of jrl-lbm_b200/csrc for the HOST
# (one loop iteration per CUDA thread) -> tests/host_emu/libjrl_lbm_emu.so.
# -ffp-contract=off mirrors the product build's -fmad=false.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
src="$here/../../jrl-lbm_b200/csrc"
if [ "${1:-}" = "reverse" ]; then  # descending serial thread order (tests/test_emu_order_independence.py)
    g++ -O2 -fPIC -shared -std=c++17 -w -ffp-contract=off -DJRL_HOST_EMU -DJRL_EMU_REVERSE \
        -I"$here" -x c++ "$src/abi.cu" "$src/ops.cu" "$src/fused.cu" "$here/emu_globals.cpp" \
        -o "$here/libjrl_lbm_emu_rev.so"
    echo "host_emu: wrote $here/libjrl_lbm_emu_rev.so"
    exit 0
fi
g++ -O2 -fopenmp -fPIC -shared -std=c++17 -w -ffp-contract=off -DJRL_HOST_EMU \
    -I"$here" -x c++ "$src/abi.cu" "$src/ops.cu" "$src/fused.cu" "$here/emu_globals.cpp" \
    -o "$here/libjrl_lbm_emu.so"
echo "host_emu: wrote $here/libjrl_lbm_emu.so"
