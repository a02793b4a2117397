"""Replaces LBM_propagation/LBM_propagation_2d.py of the reference (class at :95-186)."""
import src.LBM  # noqa: F401  (library check + sys.path)
from jrl_solver import LBMPropagation2d  # noqa: E402,F401
