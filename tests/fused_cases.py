"""Fused-step cases shared by the CPU (host-emulation) and GPU test modules: each entry runs a
scenario twice -- through the four solver objects and through ONE jrl_step_f32 launch per step --
and the results must agree bit for bit."""
import numpy as np

import jrl_scenarios as S
from util import compare_outputs


def against_oracle(oracle, fn, kw, got, case, skip=()):
    """The device path against the ORACLE on the same kwargs (the walk-loop rows must not only
    agree with the product's own host-op path): the identical scenario body runs on the numpy oracle behind the
    reference's class names (oracle/oracle_api.py), CPU tensors; integers / masks exact, floats within the bar."""
    import torch
    kw = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in kw.items() if k not in ("fused", "device_tip")}
    want = getattr(S, fn)(oracle, torch.device("cpu"), **kw)
    return compare_outputs(got, {k: v for k, v in want.items() if v is not None}, skip=skip, case=case)

FUSED_CASES = {
    "shear_bgk": ("shear_layer", dict(res=(40, 32), steps=25, operator="bgk")),
    "shear_bgk_odd": ("shear_layer", dict(res=(21, 19), steps=10, operator="bgk")),       # VEC = 1 path
    "shear_bgk_w2": ("shear_layer", dict(res=(16, 22), steps=10, operator="bgk")),        # VEC = 2 path
    "shear_mrt": ("shear_layer", dict(res=(24, 32), steps=20, operator="mrt", kbc_type=133)),
    "shear_kbc_a": ("shear_layer", dict(res=(24, 24), steps=10, operator="kbc", kbc_type=133)),
    "shear_kbc_d": ("shear_layer", dict(res=(24, 24), steps=10, operator="kbc", kbc_type=138)),
    "cylinder": ("cylinder", dict(res=(64, 48), D=8, steps=40, center=(24.0, 22.0))),
    "conc": ("jrl_conc", dict(res=(48, 32), balance_steps=15, walk_steps=40, radius_obs=2.50000000000001,
                              inflow_height=8.50000000000001, Rg=3, vel_obs_conv=0.03)),
    "conc_positive": ("jrl_conc", dict(res=(40, 28), balance_steps=5, walk_steps=20, radius_obs=2.50000000000001,
                                       inflow_height=8.50000000000001, Rg=3, vel_obs_conv=0.03, is_negative=False,
                                       clamp=False)),
}


def run_pair(api, device, name):
    fn, kw = FUSED_CASES[name]
    chain = getattr(S, fn)(api, device, **kw)
    fused = getattr(S, fn)(api, device, fused=True, **kw)
    return chain, fused


GEOMS = {"cylinder": dict(Rg=3), "plane": dict(Rg=200), "45degree": dict(Rg=3, geom="45degree", inflow_height=12.50000000000001)}


def assert_close(a, b, exact, case=""):
    """bit-identical, or (45-degree tip on the CPU build only: the scenario's torch-CPU `x / sqrt(2)` is a true division,
    the kernel follows torch-CUDA's reciprocal multiply, so phi differs in the last bit) within the usual bars."""
    if exact:
        return assert_identical(a, b)
    compare_outputs({k: v for k, v in b.items() if isinstance(v, np.ndarray)},
                    {k: v for k, v in a.items() if isinstance(v, np.ndarray)}, case=case)


def assert_identical(chain, fused):
    for k, w in chain.items():
        if isinstance(w, np.ndarray):
            assert np.array_equal(np.asarray(fused[k]), w, equal_nan=True), f"{k}: fused step differs from the operator chain"


def check_sums(api, device, res=(40, 64), force_vec=0):
    """FusedStep.step(sums=...) == sums of the stored moments over the non-OBSTACLE cells."""
    import torch
    import jrl_step
    H, W = res
    st = S.random_state(res, 21, device)
    P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=device, tau=0.7, contact_angle=None)
    fs = jrl_step.FusedStep(P, operator="bgk")
    sums = jrl_step.new_sums(device)
    f, r, u = fs.step(st["f"], st["flags"], st["rho"], st["vel"], phi_obs=st["phi"], inflow_density=1.0, sums=sums,
                      force_vec=force_vec)
    tot = jrl_step.total_sums(sums).cpu().numpy()
    fluid = (st["flags"] != 2).cpu().numpy()
    r, u = r.cpu().numpy().astype(np.float64), u.cpu().numpy().astype(np.float64)
    want = np.array([(r * fluid).sum(), (r * u[:, 0:1] * fluid).sum(), (r * u[:, 1:2] * fluid).sum(), fluid.sum()])
    assert np.allclose(tot, want, rtol=2e-6, atol=1e-6), (tot, want)


def tile_mask_reference(flags, tw):
    """numpy statement of jrl_tile_mask_u8: segment (row, c) is 1 iff every cell of [c*tw - 1, c*tw + cw]
    (periodic in x) is FLUID."""
    B, _, Hs, W = flags.shape
    tpr = (W + tw - 1) // tw
    fl = flags[:, 0] == 1
    out = np.zeros((B, Hs, tpr), np.uint8)
    for c in range(tpr):
        x0, x1 = c * tw, min(W, (c + 1) * tw)
        cols = [(x0 - 1) % W] + list(range(x0, x1)) + [x1 % W]
        out[:, :, c] = fl[:, :, cols].all(axis=2)
    return out


def check_tile_mask(api, device, res=(9, 1040)):
    assert res[0] >= 7
    """jrl_tile_mask_u8 against the numpy statement, on mostly-FLUID flags with defects placed at segment
    edges, row ends (periodic wrap) and in the partial last segment."""
    import ctypes
    import torch
    import jrl_native as N
    H, W = res
    lib = N.load()
    n = lib.jrl_tile_mask_bytes(2, H, W)
    tpr = n // (2 * H)
    tw = -(-W // tpr) if tpr > 1 else W
    # recover the library's tile width from the segment count (W chosen so that it is not ambiguous)
    for cand in (256, 512, 1024, 2048):
        if -(-W // cand) == tpr:
            tw = cand
            break
    flags = torch.ones((2, 1, H, W), dtype=torch.uint8)
    flags[0, 0, 1, 0] = 2            # first cell: also the right neighbour of the last segment
    flags[0, 0, 2, W - 1] = 8        # last cell: left neighbour of segment 0
    flags[0, 0, 3, tw - 1] = 2       # last cell of segment 0 = left neighbour of segment 1
    flags[0, 0, 4, tw] = 16          # first cell of segment 1 = right neighbour of segment 0
    flags[1, 0, 5, W - 3] = 0        # inside the partial last segment
    flags[1, 0, 6, tw // 2] = 32
    flags = flags.to(device)
    mask = torch.zeros((n,), dtype=torch.uint8, device=device)
    N.call("jrl_tile_mask_u8", ctypes.c_void_p(flags.data_ptr()), ctypes.c_void_p(mask.data_ptr()), 2, H, W, like=flags)
    got = mask.cpu().numpy().reshape(2, H, tpr)
    want = tile_mask_reference(flags.cpu().numpy(), tw)
    assert np.array_equal(got, want), (got, want)
    assert want.sum() < want.size and want.sum() > 0


RUN_CASES = {
    "shear_bgk": ("shear_layer", dict(res=(40, 32), steps=25, operator="bgk", snapshots=(10,))),
    "shear_mrt": ("shear_layer", dict(res=(24, 32), steps=20, operator="mrt", kbc_type=133)),
    "shear_kbc_a": ("shear_layer", dict(res=(24, 24), steps=9, operator="kbc", kbc_type=133)),
    "cylinder": ("cylinder", dict(res=(64, 48), D=8, steps=41, center=(24.0, 22.0), snapshots=(20,))),
}


def check_run_equals_steps(api, device, name, oracle=None):
    """jrl_run_f32 (the time loop below the ABI) == one jrl_step_f32 per step, bit for bit -- and == the oracle."""
    fn, kw = RUN_CASES[name]
    stepped = getattr(S, fn)(api, device, fused=True, **kw)
    ran = getattr(S, fn)(api, device, fused="run", **kw)
    assert_identical(stepped, ran)
    if oracle is not None:
        against_oracle(oracle, fn, kw, ran, name)


def check_run_sums_series(api, device):
    """Per-step sums series of FusedStep.run == the sums of single steps; moments-as-outputs mode."""
    import torch
    import jrl_native as N
    import jrl_step
    H, W = 24, 32
    P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=device, tau=0.7, contact_angle=None)
    fs = jrl_step.FusedStep(P, operator="bgk")
    rho, vel = S.shear_layer_state((H, W), device)
    f0 = api.SimulationRunner(P).create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
    flags = torch.ones((1, 1, H, W), dtype=torch.uint8, device=device)
    n = 7
    series = torch.zeros((n, N.SUM_SLOTS, 4), dtype=torch.float64, device=device)
    f_run, r_run, v_run = fs.run(f0.clone(), flags, None, None, n, sums=series)
    f = f0.clone()
    for i in range(n):
        one = jrl_step.new_sums(device)
        f, r, v = fs.step(f, flags, None, None, sums=one)
        assert torch.allclose(series[i].sum(0), one.sum(0), rtol=1e-11, atol=1e-9)  # (atomics: summation order varies)
    assert torch.equal(f_run, f) and torch.equal(r_run, r) and torch.equal(v_run, v)


def tip_snapshots(api, device, res, n=6, r0=2.5, inflow_height=8.50000000000001, x_interval=0.1):
    """[n,2,H,W] velocity snapshots at tip travel 0, x_interval, ... (the demos read them from the stage-A
    HDF5 file, demo_2d_LBM_JRL_C_cylinder.py:276-278); here the analytic stand-in evaluated once."""
    import torch
    H, W = res
    mesh = api.create_2d_meshgrid_tensor(size=(1, 1, H, W), device=device, dtype=torch.float32) + 0.5
    return torch.cat([S.analytic_tip_velocity(mesh, inflow_height + k * x_interval, r0) for k in range(n)], dim=0).contiguous()


def check_device_tip(api, device, is_negative=True, fused=True, oracle=None, geom="cylinder", exact=True):
    """Stage-B walk loop with the per-step host work done by ONE launch (jrl_tip_prepare_conc_f32) equals
    the demo's torch-op version bit for bit: fields, flags, flux series."""
    res = (64, 32) if geom == "45degree" else (48, 32)
    g = dict(inflow_height=8.50000000000001, **GEOMS[geom]) if "inflow_height" not in GEOMS[geom] else dict(GEOMS[geom])
    snaps = tip_snapshots(api, device, res, inflow_height=g["inflow_height"])
    kw = dict(res=res, balance_steps=5, walk_steps=45, radius_obs=2.50000000000001,
              vel_obs_conv=0.03, is_negative=is_negative, clamp=is_negative and geom != "45degree", vel_snaps=snaps,
              x_interval=0.1, fused=fused, **g)
    host = S.jrl_conc(api, device, **kw)
    dev = S.jrl_conc(api, device, device_tip=True, **kw)
    assert_close(host, dev, exact)
    assert host["walks"] == dev["walks"] and np.abs(host["current"]).max() > 0
    if fused:
        # ... and the whole walk loop in one library call (jrl_conc_walk_f32): fields and flags bit for bit; the
        # flux records come from the device strip reduction (xor tree) instead of torch.sum: 1e-5
        run = S.jrl_conc(api, device, device_tip="run", **kw)
        assert_close({k: host[k] for k in ("C", "h", "flags", "C_balance")}, run, exact)
        assert abs(run["walks"] - host["walks"]) < 1e-9
        for k in ("inflow", "current"):
            a, b = np.asarray(run[k], np.float64), np.asarray(host[k], np.float64)
            assert np.linalg.norm(a - b) <= 1e-5 * np.linalg.norm(b), k
    if oracle is not None:
        # the demo's `vel_snaps` interpolation branch (demo_2d_LBM_JRL_C_cylinder.py:294-303) on the oracle, against
        # the device-side tip motion and against the whole walk loop below the ABI
        against_oracle(oracle, "jrl_conc", kw, dev, "jrl_conc_device_tip")
        if fused:
            against_oracle(oracle, "jrl_conc", kw, run, "jrl_conc_walk")


def check_fluid_walk(api, device, gravity=2e-5, oracle=None, geom="cylinder", exact=True):
    """Stage-A loop in one library call (jrl_fluid_walk_f32) equals the demo's loop (solver objects + host torch ops
    for the tip) bit for bit."""
    g = dict(inflow_height=8.50000000000001, **GEOMS[geom]) if "inflow_height" not in GEOMS[geom] else dict(GEOMS[geom])
    kw = dict(res=(64, 32) if geom == "45degree" else (48, 32), steps=70, radius_obs=2.50000000000001, vel_obs=0.02,
              gravity_strength=gravity, snapshots=(25,), **g)
    host = S.jrl_fluid(api, device, **kw)
    run = S.jrl_fluid(api, device, device_tip="run", **kw)
    assert_close(host, run, exact, case="a_" + geom)
    assert abs(host["walks"] - run["walks"]) < 1e-12
    if oracle is not None:
        against_oracle(oracle, "jrl_fluid", kw, run, "jrl_fluid_walk")
