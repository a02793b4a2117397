"""The reference's demo scripts, BYTE-FOR-BYTE UNMODIFIED, driving the drop-in.

Each case of tests/demo_cases.py executes the reference's own script file (sha256 pinned in the golden) with
jrl-lbm_b200/ first on sys.path, so `from src.LBM.simulation import ...` binds to this repo's solver objects, and
compares main()'s captured state with tests/golden/demo_<case>.npz -- which oracle/gen_golden_demos.py wrote by running
the SAME file on the reference's own package.  Sizes and step counts are the demos' native ones: C1 / C2
130 x 130 after 100 and 1000 steps, C3 800 x 400 after 100 and 1000, and batch_validation.py's four stage-A -> stage-B
pairs at 128 x 64 (45 degree +-gravity, cylinder, infinite plane; ~2 000 + 500 + ~2 000 steps each).

  * CPU container (`-m "not gpu"`): host-emulated kernels; C1-C3 stop at the 100-step snapshot.
  * B200 (`-m gpu`): libjrl_lbm.so, full step counts.

Bars: flags / masks / step counts bit-exact; fields and record columns <= 1e-5 relative L2 (BASELINE north_star).
The scripts are read from /root/reference/demo or from baseline/_ref/demo (tools/make_baseline_ref.sh; travels to
the GPU box); without either the tests skip -- nothing under tests/golden depends on them.
"""
import os

import numpy as np
import pytest

import demo_runner as D
from conftest import GOLDEN, have_cuda
from demo_cases import DEMO_CASES
from util import field_tol, rel_l2

pytestmark = pytest.mark.skipif(D.demo_source_dir() is None, reason="reference demo scripts not available")


# The cylinder stage A has an OUTFLOW row along the bottom edge next to the moving tip, and the reference's kernel
# has a data race there (propagation.cu:195-255): on a B200 at 128 x 64 the racing reads of that row return
# a MIX of the neighbour's old and new populations in 9 launches out of 10 (tools/exp_outflow_race.py,
# profiles/r2_outflow_race.json; the right-hand OUTFLOW column, whose neighbour sits in the same warp, always reads the
# new value).  The reference's CUDA run of this case therefore differs from its own CPU run by 2e-2 and is not
# reproducible; the deterministic reading (neighbour's post-stream value = the CPU fixtures) is the pin for this pair.
# The 45-degree and infinite-plane pairs, whose flows are insensitive to that row, match the CUDA run to 1e-6.
REFCUDA_RACY = {"a_cylinder_g9", "b_cylinder_g9"}


def _golden(name, refcuda=False):
    """demo_<name>.npz: written by the reference on the CPU (oracle/gen_golden_demos.py);
    demo_refcuda_<name>.npz: written by the reference's CUDA build on a B200 (tools/bench_ref_cuda.py).  The B200
    tests prefer the latter: the 1000-step shear layers are chaotic enough that the reference's CPU and CUDA runs
    differ from EACH OTHER by 7e-2 in u at step 1000 (1.5e-6 at step 100), so only a same-arithmetic comparison
    says anything there -- and this library reproduces the CUDA run of C1 bit for bit."""
    path = os.path.join(GOLDEN, f"demo_refcuda_{name}.npz")
    if not (refcuda and os.path.exists(path)) or name in REFCUDA_RACY:
        path = os.path.join(GOLDEN, f"demo_{name}.npz")
    if not os.path.exists(path):
        pytest.skip(f"golden {path} not generated")
    with np.load(path, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def _compare(name, got, want, cpu):
    """ints bit-exact, floats rel-L2; returns the worst float error and where."""
    worst = (0.0, "")
    case = DEMO_CASES[name]
    cpu_only = cpu and case.get("cpu_steps", "all") != "all"
    for k, w in want.items():
        if cpu_only and (k.startswith("final.") or k in ("last_step",)):
            continue
        assert k in got, f"{name}: missing {k}"
        g = got[k]
        if w.dtype.kind in "iub":
            assert np.array_equal(g, w), f"{name}: {k} differs (integer / mask / step count: must be bit-exact)"
        elif w.dtype.kind in "US":
            assert str(g) == str(w), f"{name}: {k}: {g!r} vs {w!r}"
        else:
            assert g.shape == w.shape, f"{name}: {k} shape {g.shape} vs {w.shape}"
            if w.size == 0:
                continue
            e = rel_l2(g, w)
            sib = k.split(".crop_")[0] + ".stride2" if ".crop_" in k else None
            if sib in want:
                # a crop can lie where the field is still exactly zero in the reference's CPU run (C3 after 100 steps:
                # the inflow front has not arrived) while torch-CUDA's summation order -- which this library follows --
                # leaves 2e-9 of rounding residue there: measure the crop's RMS error against the field's RMS, not
                # against the crop's own (zero) norm
                rms_err = np.sqrt(np.mean((np.asarray(g, np.float64) - w) ** 2))
                e = min(e, rms_err / max(1e-30, np.sqrt(np.mean(np.asarray(want[sib], np.float64) ** 2))))
            worst = max(worst, (e, k))
            assert e <= field_tol(name, k), f"{name}: {k} relative L2 {e:.3e} > {field_tol(name, k):g}"
    return worst


@pytest.fixture(scope="module")
def workdir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("ref_demos"))


def _launches():
    import jrl_native
    return jrl_native.launch_count


# CPU container: C1-C3 (first 100 steps) and two of batch_validation's four pairs (45-degree cone, cylinder) at full
# length; with host-emulated kernels and the demos' own text dumps (1 060 Tecplot files per stage A) every further pair
# costs 1-2 minutes, and the B200 leg below runs all eleven cases.
CPU_CASES = ["c1_bgk", "c2_mrt", "c3_cylinder", "a_45degree_g0", "b_45degree_g0", "a_cylinder_g9", "b_cylinder_g9"]


@pytest.mark.parametrize("name", CPU_CASES)
def test_demo_unchanged_emu(name, emu_api, workdir):
    want = _golden(name)
    got = D.run_case(name, DEMO_CASES, workdir, cpu=True)
    worst = _compare(name, got, want, cpu=True)
    print(f"{name}: worst rel-L2 {worst[0]:.2e} ({worst[1]})")


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(DEMO_CASES))
def test_demo_unchanged_gpu(name, cuda_api, workdir):
    want = _golden(name, refcuda=True)
    n0 = _launches()
    got = D.run_case(name, DEMO_CASES, workdir, cpu=False)
    assert _launches() > n0, "the demo ran without a single libjrl_lbm launch"
    # the reference's four calls per step reach the fused kernel: ONE library launch per step (+ get_feq_ for the
    # initial state) for C1-C3 (jrl_lazy.py); stage B one per step, stage A two (fused first half + collision)
    steps = {"c1_bgk": 1000, "c2_mrt": 1000, "c3_cylinder": 1000}.get(name)
    if steps is not None:
        assert _launches() - n0 <= steps + 4, f"{name}: {_launches() - n0} library launches for {steps} steps"
    worst = _compare(name, got, want, cpu=False)
    print(f"{name}: worst rel-L2 {worst[0]:.2e} ({worst[1]}), {_launches() - n0} library launches")
