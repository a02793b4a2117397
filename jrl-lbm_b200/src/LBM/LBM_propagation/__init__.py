from .LBM_propagation_2d import LBMPropagation2d  # noqa: F401
