"""The fused step behind the reference's four calls (jrl_lazy.py).

The same demo loop bodies (tests/jrl_scenarios.py: propagation -> [solve_boundary] -> macro_compute(_C) -> collision,
exactly the reference's call sequence) run twice -- every call launching immediately (JRL_LBM_FUSE=0 behaviour) and
with the deferred chain -- and must agree BIT FOR BIT, while the deferred run issues ONE library launch per step.
Also: anything that looks at a deferred field gets the eager result; a mutated input raises.
CPU container: host-emulated kernels; `-m gpu`: libjrl_lbm.so."""
import numpy as np
import pytest
import torch

import jrl_lazy as L
import jrl_native
import jrl_scenarios as S
from conftest import have_cuda

JRL = dict(radius_obs=2.50000000000001, inflow_height=8.50000000000001, Rg=3)
# name -> (scenario, kwargs, steps, library launches per step when fused)
CASES = {
    "c1_bgk": ("shear_layer", dict(res=(40, 32), steps=12, operator="bgk"), 12, 1),
    "c2_mrt": ("shear_layer", dict(res=(40, 32), steps=12, operator="mrt", kbc_type=133), 12, 1),
    "c2_kbc_a": ("shear_layer", dict(res=(32, 32), steps=8, operator="kbc", kbc_type=133), 8, 1),
    "c3_cylinder": ("cylinder", dict(res=(96, 48), D=8, steps=20, center=(24.0, 30.0)), 20, 1),
    "c4_conc": ("jrl_conc", dict(res=(48, 32), balance_steps=10, walk_steps=20, vel_obs_conv=0.03, **JRL), 30, 1),
    "c4_conc_noclamp": ("jrl_conc", dict(res=(40, 28), balance_steps=5, walk_steps=10, vel_obs_conv=0.03, is_negative=False,
                                         clamp=False, **JRL), 15, 1),
}


def _run(api, dev, name, enabled):
    fn, kw, steps, _ = CASES[name]
    old = L.ENABLED
    L.ENABLED = enabled
    try:
        n0, f0 = jrl_native.launch_count, L.stats["fused"]
        out = getattr(S, fn)(api, dev, **kw)
        return out, jrl_native.launch_count - n0, L.stats["fused"] - f0
    finally:
        L.ENABLED = old


def _identical(a, b):
    for k, v in a.items():
        if isinstance(v, np.ndarray):
            assert np.array_equal(v, b[k], equal_nan=True), f"{k}: deferred chain differs from the immediate launches"


def _check(api, dev, name):
    eager, n_eager, _ = _run(api, dev, name, False)
    lazy, n_lazy, fused = _run(api, dev, name, True)
    _identical(eager, lazy)
    steps, per_step = CASES[name][2], CASES[name][3]
    assert fused == steps, f"{fused} of {steps} steps took the fused launch: {L.stats['reasons']}"
    assert n_lazy < n_eager
    return n_eager, n_lazy


@pytest.mark.parametrize("name", list(CASES))
def test_deferred_equals_immediate_emu(emu_api, name):
    print(name, "launches immediate / deferred:", _check(emu_api, torch.device("cpu"), name))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_deferred_equals_immediate_gpu(cuda_api, name):
    print(name, "launches immediate / deferred:", _check(cuda_api, torch.device("cuda:0"), name))


def _stage_a(api, dev):
    kw = dict(res=(48, 32), steps=15, vel_obs=0.02, gravity_strength=2e-5, **JRL)
    L.ENABLED = False
    try:
        n0 = jrl_native.launch_count
        a = S.jrl_fluid(api, dev, **kw)
        n_eager = jrl_native.launch_count - n0
    finally:
        L.ENABLED = True
    n0 = jrl_native.launch_count
    b = S.jrl_fluid(api, dev, **kw)
    n_lazy = jrl_native.launch_count - n0
    _identical(a, b)
    assert n_eager - n_lazy == 2 * 15, (n_eager, n_lazy)   # 4 launches per step -> 2


def test_stage_a_is_two_launches_emu(emu_api):
    """The axisymmetric fluid collision needs the neighbours' new moments: streaming + wall + moments fuse into one
    launch (JRL_OP_NONE), the collision is the second; same bits as the four launches."""
    _stage_a(emu_api, torch.device("cpu"))


@pytest.mark.gpu
def test_stage_a_is_two_launches_gpu(cuda_api):
    _stage_a(cuda_api, torch.device("cuda:0"))


def _touch(api, dev):
    """Looking at a deferred field mid-step gives what the immediate launches give, and the step still completes."""
    def body(peek):
        P = api.SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, 24, 20), dt=1.0,
                                     contact_angle=None, Q=9, tau=0.7)
        R = api.SimulationRunner(parameters=P)
        prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
        s = S.random_state((24, 20), 3, dev)
        f, flags, rho, vel = s["f"], s["flags"], s["rho"], s["vel"]
        seen = []
        for step in range(4):
            r, v, f = prop.propagation(f=f, flags=flags, rho=rho, vel=vel, inflow_density=1.0,
                                       inflow_vel=torch.tensor([0.01, 0.02], device=dev))
            if peek == "f":
                seen.append(f[0, 3, 5, 5].item())
            if peek == "rho_in":
                seen.append(float(r.sum()))
            rho, vel = macro.macro_compute(dx=1.0, dt=1.0, f=f, rho=r, vel=v, flags=flags)
            if peek == "vel":
                seen.append(float((vel * vel).sum()))
            f = col.collision(dx=1.0, dt=1.0, f=f, rho=rho, vel=vel, flags=flags)
        return f.cpu().numpy(), rho.cpu().numpy(), vel.cpu().numpy(), seen
    base = body(None)
    for peek in ("f", "rho_in", "vel"):
        got = body(peek)
        for a, b in zip(base[:3], got[:3]):
            assert np.array_equal(a, b, equal_nan=True), peek
        L.ENABLED = False
        try:
            ref = body(peek)
        finally:
            L.ENABLED = True
        assert ref[3] == got[3], f"peeking at {peek}: deferred field showed different values than the immediate launch"


def test_touching_a_deferred_field_emu(emu_api):
    _touch(emu_api, torch.device("cpu"))


@pytest.mark.gpu
def test_touching_a_deferred_field_gpu(cuda_api):
    _touch(cuda_api, torch.device("cuda:0"))


def test_mutated_input_raises_emu(emu_api):
    api, dev = emu_api, torch.device("cpu")
    P = api.SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, 16, 16), dt=1.0,
                                 contact_angle=None, Q=9, tau=0.7)
    R = api.SimulationRunner(parameters=P)
    prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
    s = S.random_state((16, 16), 5, dev, io_cells=False)
    f2 = prop.propagation(f=s["f"], flags=s["flags"])
    rho, vel = macro.macro_compute(dx=1.0, dt=1.0, f=f2, rho=s["rho"], vel=s["vel"], flags=s["flags"])
    s["flags"][0, 0, 3, 3] = 2   # the recorded step can no longer be replayed as the immediate launch would have run
    with pytest.raises(RuntimeError, match="modified in place"):
        col.collision(dx=1.0, dt=1.0, f=f2, rho=rho, vel=vel, flags=s["flags"])


def test_flags_cache_is_per_tensor_object_emu(emu_api):
    """caches keyed on the flags address survive a free + reallocation at the same address."""
    import jrl_step
    api, dev = emu_api, torch.device("cpu")
    P = api.SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, 8, 512), dt=1.0,
                                 contact_angle=None, Q=9, tau=0.7)
    fs = jrl_step.FusedStep(P, operator="bgk")
    a = torch.ones((1, 1, 8, 512), dtype=torch.uint8)
    m1 = fs._tile_mask(a, 1, 8, 512)
    assert m1 is not None and fs._tile_mask(a, 1, 8, 512) is m1          # cached for the same object
    b = torch.ones((1, 1, 8, 512), dtype=torch.uint8)
    b[0, 0, 2, 40] = 2
    fs._masks[id(b)] = fs._masks[id(a)]                                   # what an address-keyed cache would do
    m2 = fs._tile_mask(b, 1, 8, 512)
    assert not torch.equal(m2, torch.ones_like(m2)) or m2 is not m1


def _original_tensor_still_mutated(api, dev):
    """A caller that ignores the tensors propagation() returns and keeps using the ORIGINAL rho / vel objects still
    sees streaming's in-place writes at INFLOW cells after the fused step (propagation.cu:300, 320-321)."""
    P = api.SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, 24, 20), dt=1.0,
                                 contact_angle=None, Q=9, tau=0.7)
    R = api.SimulationRunner(parameters=P)
    prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
    s = S.random_state((24, 20), 3, dev)
    outs = []
    for enabled in (False, True):
        L.ENABLED = enabled
        try:
            rho, vel = s["rho"].clone(), s["vel"].clone()
            res = prop.propagation(f=s["f"], flags=s["flags"], rho=rho, vel=vel, inflow_density=1.0,
                                   inflow_vel=torch.tensor([0.01, 0.02], device=dev))
            f2 = res[2]
            del res
            r2, v2 = macro.macro_compute(dx=1.0, dt=1.0, f=f2, rho=rho, vel=vel, flags=s["flags"])
            f3 = col.collision(dx=1.0, dt=1.0, f=f2, rho=r2, vel=v2, flags=s["flags"])
            outs.append([t.cpu().numpy().copy() for t in (rho, vel, f3)])
        finally:
            L.ENABLED = True
    for a, b in zip(*outs):
        assert np.array_equal(a, b, equal_nan=True)
    assert not np.array_equal(outs[1][0], s["rho"].cpu().numpy())   # the inflow density really was written


def test_original_tensor_still_mutated_emu(emu_api):
    _original_tensor_still_mutated(emu_api, torch.device("cpu"))


@pytest.mark.gpu
def test_original_tensor_still_mutated_gpu(cuda_api):
    _original_tensor_still_mutated(cuda_api, torch.device("cuda:0"))
