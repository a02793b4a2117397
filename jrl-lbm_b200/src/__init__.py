"""Drop-in for the reference's top-level `src` package: demo scripts do `sys.path.append("../")`
and `from src.LBM.simulation import ...` (demo_2d_LBM.py:10-15); put jrl-lbm_b200/ on sys.path
instead and the same imports resolve to the B200 library."""
