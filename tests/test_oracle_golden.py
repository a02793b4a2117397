"""CPU: pin the oracle.  The numpy restatement (oracle/lbm_oracle.py, driven through the reference's
solver-object API by oracle/oracle_api.py) must reproduce every golden fixture in tests/golden/,
which were written by the REFERENCE's own unmodified Python classes run on the CPU
(oracle/gen_golden.py -> oracle/ref_cpu.import_reference()).  Integer / mask outputs bit-exact,
float fields within 1e-5 relative L2 (util.REL_L2_TOL) -- in practice most agree to the last bit.
Also pins oracle/ref_torch_port.py (the CPU baseline loop bench.py times) to the same fixture."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S
from golden_cases import CASES
from util import compare_outputs, load_golden, rel_l2


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_reproduces_reference_golden(oracle_api, name):
    fn, kw = CASES[name]
    got = getattr(S, fn)(oracle_api, torch.device("cpu"), **kw)
    want = load_golden(name)
    assert len(want) >= 3
    compare_outputs(got, want, case=name)


def test_golden_fixtures_are_not_trivial():
    """Every fixture holds evolved, finite, non-constant fields (guards against a generator that
    silently wrote initial states)."""
    for name in CASES:
        z = load_golden(name)
        floats = [k for k, v in z.items() if v.dtype.kind == "f" and v.ndim >= 3]
        assert floats, name
        assert any(np.nanstd(z[k].astype(np.float64)) > 0 for k in floats), name
    z = load_golden("c1_shear_bgk")
    assert rel_l2(z["f"], z["f0"]) > 1e-3  # the state moved


def test_cpu_baseline_loop_matches_reference_golden():
    """oracle/ref_torch_port.ReferenceCpuLoop (reference K1 body compiled for the host + the
    reference's torch-op chain) is what bench.py times as cpu_baseline / --impl reference: it must
    compute the same C1 run as the reference's classes did."""
    import ref_cpu
    if not ref_cpu.have_ref_lib():
        pytest.skip("oracle/_ref/libjrl_ref.so not built (needs /root/reference; prebuilt on the GPU box)")
    import ref_torch_port as RP
    fn, kw = CASES["c1_shear_bgk"]
    z = load_golden("c1_shear_bgk")
    H, W = kw["res"]
    rho, vel = S.shear_layer_state((H, W), torch.device("cpu"))
    tau = float(z["tau"]) if "tau" in z else None
    if tau is None:
        pytest.skip("fixture carries no tau")
    loop = RP.ReferenceCpuLoop(tau)
    f = RP.get_feq(1.0, 1.0, rho, vel)
    assert rel_l2(f.numpy(), z["f0"]) <= 1e-6
    flags = torch.ones((1, 1, H, W), dtype=torch.uint8)
    for _ in range(kw["steps"]):
        f, rho, vel = loop.step(f, rho, vel, flags)
    for k, t in (("f", f), ("rho", rho), ("vel", vel)):
        assert rel_l2(t.numpy(), z[k]) <= 1e-5, k
