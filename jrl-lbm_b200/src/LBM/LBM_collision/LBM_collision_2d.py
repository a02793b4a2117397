"""Replaces LBM_collision/LBM_collision_2d.py of the reference (class at :125-555)."""
import src.LBM  # noqa: F401
from jrl_solver import LBMCollision2d  # noqa: E402,F401
