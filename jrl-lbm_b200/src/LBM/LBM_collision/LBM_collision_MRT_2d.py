"""Replaces LBM_collision/LBM_collision_MRT_2d.py of the reference (class at :11-184)."""
import src.LBM  # noqa: F401
from jrl_solver import LBMCollisionMRT2d  # noqa: E402,F401
