"""Replaces LBM_collision/LBM_collision_KBC_2d.py of the reference (class at :11-317)."""
import src.LBM  # noqa: F401
from jrl_solver import LBMCollisionKBC2d  # noqa: E402,F401
