"""Replaces LBM_macro_compute/LBM_macro_compute_2d.py of the reference (class at :105-361)."""
import src.LBM  # noqa: F401
from jrl_solver import LBMMacroCompute2d  # noqa: E402,F401
