from .LBM_macro_compute_2d import LBMMacroCompute2d  # noqa: F401
