"""Shared helpers of the test-suite."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# BASELINE.json north_star: "macroscopic fields and obstacle forces within 1e-5 relative L2 after N
# steps in fp32"; masks and indexing bit-exact.
REL_L2_TOL = 1e-5


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    fin = np.isfinite(a) & np.isfinite(b)
    assert np.array_equal(np.isfinite(a), np.isfinite(b)), "non-finite patterns differ"
    assert np.array_equal(a[~fin], b[~fin], equal_nan=True), "inf/nan values differ"
    return float(np.linalg.norm(a[fin] - b[fin]) / max(1e-30, np.linalg.norm(b[fin])))


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def compare_outputs(got, want, tol=REL_L2_TOL, skip=()):
    """Every key of `want`: integer arrays bit-exact, float arrays within `tol` relative L2,
    python scalars to 1e-12.  Returns {key: rel_l2} for reporting."""
    report = {}
    for k, w in want.items():
        if k in skip:
            continue
        assert k in got, f"missing output {k}"
        g = got[k]
        if g is None:
            continue
        w = np.asarray(w)
        g = np.asarray(g)
        if w.dtype.kind in "iub":
            assert np.array_equal(g, w), f"{k}: integer / mask output differs"
        elif w.ndim == 0:
            assert abs(float(g) - float(w)) <= 1e-12 * max(1.0, abs(float(w))), f"{k}: {g} vs {w}"
        else:
            assert g.shape == w.shape, f"{k}: shape {g.shape} vs {w.shape}"
            e = rel_l2(g, w)
            report[k] = e
            assert e <= tol, f"{k}: relative L2 {e:.3e} > {tol:g}"
    return report
