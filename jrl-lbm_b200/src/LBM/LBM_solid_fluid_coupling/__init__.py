from .LBM_boundary_2d import LBMBoundary2d  # noqa: F401
