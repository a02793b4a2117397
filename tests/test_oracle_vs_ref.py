"""CPU: pin the oracle's kernel restatements against the REFERENCE's own kernel bodies.

oracle/_ref/libjrl_ref.so holds kernel_prop_2d_forward, kernel_get_grad/div_2d_forward,
kernel_macompute(_C)_2d_forward and the solid-fluid kernel_2d_forward exactly as written in the
reference's .cu files, compiled for the host by oracle/build_ref.sh (one loop iteration per CUDA
thread).  On seeded random states holding every CellType, with and without symmetry axes, batch 1
and 2, the numpy oracle must reproduce them BIT FOR BIT (NaN patterns included).  Skipped where
the .so was never built (it needs /root/reference; the GPU box receives it prebuilt)."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S
import lbm_oracle as O
import ref_cpu

pytestmark = pytest.mark.skipif(not ref_cpu.have_ref_lib(), reason="oracle/_ref/libjrl_ref.so not built")

STATES = [((20, 24), 1, 0, 1), ((18, 14), 2, 1, 1), ((14, 18), 3, 2, 1), ((12, 16), 4, 1, 2), ((33, 47), 5, 0, 1),
          ((16, 64), 6, 2, 2)]


@pytest.fixture(scope="module")
def ker():
    return ref_cpu.RefKernels()


def _state(res, seed, axisym, B):
    st = S.random_state(res, seed, torch.device("cpu"), axisym, B=B)
    return {k: v.numpy() for k, v in st.items()}


def _same(a, b, what):
    assert a.dtype == b.dtype and a.shape == b.shape, what
    assert np.array_equal(a, b, equal_nan=True), f"{what}: oracle differs from the reference kernel body " \
        f"({int((~((a == b) | (np.isnan(a) & np.isnan(b)))).sum())} elements)"


@pytest.mark.parametrize("res,seed,axisym,B", STATES)
def test_stream_K1(ker, res, seed, axisym, B):
    s = _state(res, seed, axisym, B)
    iv = np.array([0.01, 0.05], np.float32)
    variants = [dict(), dict(phi_obs=s["phi"]), dict(inflow_vel=iv, inflow_density=1.0),
                dict(phi_obs=s["phi"], inflow_density=0.0), dict(inflow_vel=iv)]
    for kw in variants:
        r1, v1, r2, v2 = s["rho"].copy(), s["vel"].copy(), s["rho"].copy(), s["vel"].copy()
        want = ker.prop(s["flags"], s["f"], rho=r1, vel=v1, axisym=axisym, **kw)
        got = O.stream(s["flags"], s["f"], rho=r2, vel=v2, axisym=axisym, **kw)
        _same(got, want, f"K1 {sorted(kw)}")
        _same(r2, r1, "K1 rho mutation")
        _same(v2, v1, "K1 vel mutation")


@pytest.mark.parametrize("res,seed,axisym,B", STATES)
def test_grad_div_K2_K3(ker, res, seed, axisym, B):
    s = _state(res, seed, axisym, B)
    _same(O.grad(s["rho"], s["flags"], 1.0, axisym), ker.grad(s["rho"], s["flags"], 1.0, axisym), "K2")
    _same(O.grad(s["rho"], s["flags"], 0.5, axisym), ker.grad(s["rho"], s["flags"], 0.5, axisym), "K2 dx=0.5")
    _same(O.div(s["vel"], s["flags"], 1.0, axisym), ker.div(s["vel"], s["flags"], 1.0, axisym), "K3")


@pytest.mark.parametrize("res,seed,axisym,B", STATES)
def test_moments_K5_K6(ker, res, seed, axisym, B):
    s = _state(res, seed, axisym, B)
    H, W = res
    mesh = O.meshgrid(B, H, W) + np.float32(0.5)
    want = ker.macro(1.0, 1.0, s["f"], s["rho"], s["vel"], s["flags"], axisym=axisym)
    got = O.macro_compute_kernel(1.0, 1.0, s["f"], s["rho"], s["vel"], s["flags"])
    _same(got[0], want[0], "K6 rho")
    _same(got[1], want[1], "K6 vel")
    wantC = ker.macro_C(1.0, 1.0, s["f"], s["rho"], s["flags"], s["vel"], mesh, axisym)
    gotC = O.macro_compute_C_kernel(1.0, 1.0, s["f"], s["rho"], s["flags"], s["vel"], mesh, axisym)
    _same(gotC, wantC, "K5")


@pytest.mark.parametrize("res,seed,axisym,B", STATES)
def test_wall_K7(ker, res, seed, axisym, B):
    s = _state(res, seed, axisym, B)
    for obs_vel in (s["obs_vel"], None):
        want = ker.wall(1.0, 1.0, 0.8, s["rho"], s["vel"], s["flags"], s["f"], s["phi"], obs_vel, axisym=axisym)
        got = O.wall(1.0, 1.0, 0.8, s["vel"], s["flags"], s["f"], s["phi"], obs_vel, axisym)
        _same(got, want, "K7")
