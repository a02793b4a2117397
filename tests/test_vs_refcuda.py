"""Parity against the reference AS ITS USERS RUN IT: tests/golden/refcuda_*.npz are the
outputs of the reference's own fast-math CUDA build on a B200 (tools/bench_ref_cuda.py --role ref), for every case of
tests/golden_cases.py.  tests/refcuda_cases.py states which outputs must agree to the last bit and why the others
only agree to 1e-5.

  * `-m gpu`: libjrl_lbm.so on the B200 vs those fixtures -- bit-exactness asserted key by key.
  * CPU container: the numpy oracle and the host-emulated kernels vs the same fixtures (the oracle restates torch's
    CUDA arithmetic, so the bit-exact set holds there too, except where a 9x9 matrix inverse is involved).
"""
import os

import numpy as np
import pytest
import torch

import jrl_scenarios as S
import refcuda_cases as R
from conftest import GOLDEN
from golden_cases import CASES
from util import REL_L2_TOL, field_tol, rel_l2


def _load(name):
    path = os.path.join(GOLDEN, f"refcuda_{name}.npz")
    if not os.path.exists(path):
        pytest.skip(f"{path} missing (written on a GPU box by tools/bench_ref_cuda.py)")
    with np.load(path, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def _bits_differ(a, b):
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return int((a.view(np.uint32) != b.view(np.uint32)).sum())


def check_case(name, got, want, exact_keys, exact_scenarios):
    """Returns {key: (rel_l2, differing bits)}; asserts the tables of refcuda_cases.py."""
    report = {}
    flags = want.get("inputs_flags")
    batch_bug = name == "ops_batch2"
    for k, w in want.items():
        if w.ndim == 0:
            continue
        g = np.asarray(got[k])
        if w.dtype.kind in "iub":
            assert np.array_equal(g, w), f"{name}: {k} integer / mask output differs"
            continue
        if k in R.STREAM_KEYS and flags is not None:       # the reference's OUTFLOW race
            keep = np.broadcast_to(flags != R.OUTFLOW, w.shape)
            g, w = np.where(keep, g, 0), np.where(keep, w, 0)
        if batch_bug and k in R.BATCH_BUG_KEYS:            # reference bug for B > 1 (see refcuda_cases.py)
            g, w = g[0:1], w[0:1]
        e, nb = rel_l2(g, w), _bits_differ(g, w)
        report[k] = (e, nb)
        assert e <= field_tol(name, k), f"{name}: {k} relative L2 {e:.3e} vs the reference-as-run"
        must = k in exact_keys if name.startswith("ops_") else (
            name in exact_scenarios and (exact_scenarios[name] is None or k in exact_scenarios[name]))
        if must:
            assert nb == 0, f"{name}: {k}: {nb} of {w.size} values differ in the last bits from the reference-as-run"
    return report


def _fmt(report):
    exact = sorted(k for k, (e, nb) in report.items() if nb == 0)
    rest = {k: f"{e:.1e}/{nb}" for k, (e, nb) in report.items() if nb}
    return f"bit-exact: {exact}; others (rel-L2/bits): {rest}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_cuda_vs_reference_as_run(cuda_api, name):
    fn, kw = CASES[name]
    want = _load(name)
    got = getattr(S, fn)(cuda_api, torch.device("cuda:0"), **kw)
    print(name, _fmt(check_case(name, got, want, R.GPU_EXACT_KEYS, R.GPU_EXACT_SCENARIOS)))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1_shear_bgk", "c2_shear_mrt", "c3_cylinder", "c4_jrl_conc"])
def test_fused_vs_reference_as_run(cuda_api, name):
    fn, kw = CASES[name]
    want = {k: v for k, v in _load(name).items() if not k.startswith("snap_")}
    got = getattr(S, fn)(cuda_api, torch.device("cuda:0"), fused=True, **kw)
    print(name, _fmt(check_case(name, got, want, R.GPU_EXACT_KEYS, R.GPU_EXACT_SCENARIOS)))


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_vs_reference_as_run(oracle_api, name):
    fn, kw = CASES[name]
    want = _load(name)
    got = getattr(S, fn)(oracle_api, torch.device("cpu"), **kw)
    print(name, _fmt(check_case(name, got, want, R.EXACT_KEYS, R.EXACT_SCENARIOS)))


@pytest.mark.parametrize("name", ["ops_plain", "ops_axis_x", "ops_axis_y", "c1_shear_bgk", "c4_jrl_conc"])
def test_emu_vs_reference_as_run(emu_api, name):
    fn, kw = CASES[name]
    want = _load(name)
    got = getattr(S, fn)(emu_api, torch.device("cpu"), **kw)
    print(name, _fmt(check_case(name, got, want, R.EXACT_KEYS, R.EXACT_SCENARIOS)))
