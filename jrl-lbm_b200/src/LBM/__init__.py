"""`src.LBM` -- same sub-packages as the reference (simulation, utils, LBM_propagation,
LBM_collision, LBM_macro_compute, LBM_solid_fluid_coupling).  The reference JIT-compiles four
CUDA extensions at import (LBM_propagation_2d.py:11-20 and siblings); this package opens ONE
prebuilt C-ABI library, libjrl_lbm.so, and fails loudly if it is missing."""
import os
import sys

_pkg_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _pkg_root not in sys.path:
    sys.path.insert(0, _pkg_root)  # jrl_native / jrl_solver live next to `src/`

import jrl_native as _native  # noqa: E402

_native.load()
