"""Replaces LBM_solid_fluid_coupling/LBM_boundary_2d.py of the reference (class at :76-158)."""
import src.LBM  # noqa: F401
from jrl_solver import LBMBoundary2d  # noqa: E402,F401
