"""Randomised differential test of the deferred chain (jrl_lazy.py) against USER CODE the demos do not contain.

A reference user may do anything with the tensors the four calls return -- read them, clone them, slice them, write
into them, hand them to other torch functions, keep old ones, skip a call -- and the drop-in has to give what the
immediate launches give, whatever is done and in whatever order.  Each "program" below is a seeded random sequence of
such actions inserted between the calls of a few time steps (two solver sets interleaved in some of them, as the coupled
demos do); it runs once with every call launching immediately (JRL_LBM_FUSE=0 behaviour) and once deferred.  Bars:

  * every observed value and the final fields are BIT-IDENTICAL, or
  * the deferred run raises the documented RuntimeError ("modified in place") -- accepted only for programs that
    really wrote into a recorded input before the step was launched; never a silently different result.

Host-emulated kernels (CPU container); the GPU suite covers the same layer through test_lazy_fusion.py and the
unmodified demos."""
import random

import numpy as np
import pytest
import torch

import jrl_lazy as L
import jrl_scenarios as S

DEV = torch.device("cpu")
H, W = 18, 14

# actions after propagation (A), after macro_compute (B), after collision (C)
# ("none" and the metadata reads keep the step fusable; the writes into recorded inputs are rare on purpose: most programs
# must reach the bit-identical bar, not the documented error)
ACTIONS_A = ["none"] * 20 + ["meta_f"] * 3 + ["read_f", "clone_f", "inplace_f", "read_rho_in", "cpu_f", "slice_f", "cat_f",
                                              "drop_f_keep_going"] * 2 + ["write_flags", "write_rho_orig"]
ACTIONS_B = ["none"] * 20 + ["meta_rho"] * 3 + ["read_vel", "inplace_rho", "rebind_vel", "slice_vel", "norm_vel", "detach_rho",
                                                "index_put_vel"] * 2 + ["write_flags"]
ACTIONS_C = ["none"] * 4 + ["read_f", "read_old_rho", "inplace_f", "clone_all", "read_stale_stream"]
MUTATING = {"write_flags", "write_rho_orig"}


def _solvers(api, tau):
    P = api.SimulationParameters(dim=2, dtype=torch.float32, device=DEV, simulation_size=(1, 1, H, W), dt=1.0,
                                 contact_angle=None, Q=9, tau=tau)
    R = api.SimulationRunner(parameters=P)
    return R.create_propagation(), R.create_macro_compute(), R.create_collision()


def _program(api, seed, enabled):
    """Run program `seed`; returns (observations, final arrays, set of mutating actions used)."""
    rng = random.Random(seed)
    L.ENABLED = enabled
    used = set()
    try:
        two = rng.random() < 0.3
        sets = [_solvers(api, 0.7)] + ([_solvers(api, 0.9)] if two else [])
        states = []
        for k in range(len(sets)):
            s = S.random_state((H, W), 11 + seed % 7 + k, DEV, io_cells=rng.random() < 0.7)
            states.append(dict(f=s["f"], flags=s["flags"], rho=s["rho"], vel=s["vel"]))
        inflow = rng.random() < 0.6
        seen = []

        def obs(x):
            seen.append(np.asarray(x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x).copy())

        for step in range(rng.randint(2, 4)):
            order = list(range(len(sets)))
            rng.shuffle(order)
            mid = {}
            # first halves of every solver set, then second halves: interleaved chains
            for k in order:
                prop, macro, col = sets[k]
                st = states[k]
                old_rho = st["rho"]
                kw = dict(inflow_density=1.0, inflow_vel=torch.tensor([0.01, 0.02])) if inflow else {}
                res = prop.propagation(f=st["f"], flags=st["flags"], rho=st["rho"], vel=st["vel"], **kw)
                r_in, v_in, f = res if isinstance(res, (tuple, list)) else (st["rho"], st["vel"], res)
                a = rng.choice(ACTIONS_A)
                used.add(a)
                if a == "read_f":
                    obs(f.sum())
                elif a == "clone_f":
                    f = f.clone()
                elif a == "inplace_f":
                    f[:, 1, 2, 3] += 0.001
                elif a == "read_rho_in":
                    obs(r_in[0, 0, 1])
                elif a == "cpu_f":
                    obs(f.cpu()[0, 4])
                elif a == "slice_f":
                    obs(f[:, 2:4, 1:5, ::3])
                elif a == "cat_f":
                    obs(torch.cat([f, f], dim=1).mean())
                elif a == "meta_f":
                    assert tuple(f.shape) == (1, 9, H, W) and f.dtype == torch.float32 and f.dim() == 4 and f.numel() == 9 * H * W
                elif a == "write_flags":
                    st["flags"][0, 0, 5, 5] = 2
                elif a == "write_rho_orig":
                    st["rho"][0, 0, 4, 4] = 1.25
                elif a == "drop_f_keep_going":
                    # the user throws this step's streaming away and streams again
                    res = prop.propagation(f=st["f"], flags=st["flags"], rho=st["rho"], vel=st["vel"], **kw)
                    r_in, v_in, f = res if isinstance(res, (tuple, list)) else (st["rho"], st["vel"], res)
                mid[k] = (f, r_in, v_in, old_rho)
            for k in order:
                prop, macro, col = sets[k]
                st = states[k]
                f, r_in, v_in, old_rho = mid[k]
                rho, vel = macro.macro_compute(dx=1.0, dt=1.0, f=f, rho=r_in, vel=v_in, flags=st["flags"])
                b = rng.choice(ACTIONS_B)
                used.add(b)
                view = None
                if b == "read_vel":
                    obs((vel * vel).sum())
                elif b == "inplace_rho":
                    rho.mul_(1.0001)
                elif b == "rebind_vel":
                    vel = vel * 1.0
                elif b == "slice_vel":
                    view = vel[:, 0:1]
                elif b == "norm_vel":
                    obs(torch.linalg.vector_norm(vel, dim=1).max())
                elif b == "meta_rho":
                    assert rho.size(1) == 1 and rho.device == DEV and rho.is_contiguous()
                elif b == "detach_rho":
                    rho = rho.detach()
                elif b == "write_flags":
                    st["flags"][0, 0, 6, 6] = 2
                elif b == "index_put_vel":
                    vel[:, :, 0, :] = 0.0
                f_new = col.collision(dx=1.0, dt=1.0, f=f, rho=rho, vel=vel, flags=st["flags"])
                if view is not None:
                    obs(view)
                c = rng.choice(ACTIONS_C)
                used.add(c)
                if c == "read_f":
                    obs(f_new[0, :, 3, 3])
                elif c == "read_old_rho":
                    obs(old_rho)            # pre-step rho object: streaming's in-place INFLOW writes must be there
                elif c == "inplace_f":
                    f_new[:, 0] *= 1.0001
                elif c == "clone_all":
                    f_new, rho, vel = f_new.clone(), rho.clone(), vel.clone()
                elif c == "read_stale_stream":
                    obs(f.mean())           # the streamed populations, looked at AFTER the fused step ran
                st.update(f=f_new, rho=rho, vel=vel)
        final = [np.asarray(st[k].detach().cpu().numpy()).copy() for st in states for k in ("f", "rho", "vel", "flags")]
        return seen, final, used
    finally:
        L.ENABLED = True


@pytest.mark.parametrize("block", range(6))
def test_random_user_programs_emu(emu_api, block):
    fused0 = L.stats["fused"]
    raised = 0
    for seed in range(block * 20, block * 20 + 20):
        want_seen, want_final, used = _program(emu_api, seed, enabled=False)
        try:
            got_seen, got_final, _ = _program(emu_api, seed, enabled=True)
        except RuntimeError as e:
            assert "modified in place" in str(e), f"program {seed}: {e}"
            assert used & MUTATING, f"program {seed} raised without writing into a recorded input: {sorted(used)}"
            raised += 1
            continue
        assert len(want_seen) == len(got_seen), f"program {seed}"
        for i, (a, b) in enumerate(zip(want_seen, got_seen)):
            assert np.array_equal(a, b, equal_nan=True), f"program {seed}: observation {i} differs ({sorted(used)})"
        for i, (a, b) in enumerate(zip(want_final, got_final)):
            assert np.array_equal(a, b, equal_nan=True), f"program {seed}: final field {i} differs ({sorted(used)})"
    print(f"block {block}: 20 programs, {raised} raised (documented), {L.stats['fused'] - fused0} steps fused")
    assert raised <= 10 and L.stats["fused"] - fused0 >= 8, "the programs no longer exercise both the fused and the replayed path"


CLAMPS = {
    "torch.clamp(C, 0, 1)": lambda C: torch.clamp(C, 0, 1),          # the demos' spelling: fused into the step
    "C.clamp(0, 1)": lambda C: C.clamp(0, 1),
    "C.clamp_(0, 1)": lambda C: C.clamp_(0, 1),
    "torch.clamp(C, min=0, max=1)": lambda C: torch.clamp(C, min=0, max=1),
    "torch.clamp(C, 0.0, 1.0)": lambda C: torch.clamp(C, 0.0, 1.0),
    "clamp(0, 2)": lambda C: torch.clamp(C, 0, 2),                   # a different bound: must not be taken for the fused clamp
    "clip": lambda C: torch.clip(C, 0, 1),
    "where": lambda C: torch.where(C > 1, torch.ones_like(C), C),
    "none": lambda C: C,
}


@pytest.mark.parametrize("spelling", list(CLAMPS))
@pytest.mark.parametrize("axisym", [0, 1])
def test_concentration_chain_clamp_spellings_emu(emu_api, spelling, axisym):
    """The concentration demos clamp C between macro_compute_C and collision (demo_2d_LBM_JRL_C_cylinder.py:324); only
    the exact call the fused kernel implements may be absorbed, every other spelling has to give the immediate result."""
    api = emu_api
    outs = []
    for enabled in (False, True):
        L.ENABLED = enabled
        try:
            P = api.SimulationParameters(dim=2, dtype=torch.float32, device=DEV, simulation_size=(1, 1, H, W), dt=1.0,
                                         contact_angle=None, Q=9, tau=0.8, axisymmetric_type=axisym)
            R = api.SimulationRunner(parameters=P)
            prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
            col.preset_KBC(dx=1.0, dt=1.0, tau=0.8, tau_D=0.9)
            s = S.random_state((H, W), 4, DEV, axisym)
            mesh = api.create_2d_meshgrid_tensor(size=(1, 1, H, W), device=DEV, dtype=torch.float32) + 0.5
            h, C, flags, vel = s["f"], s["rho"] * 1.2, s["flags"], s["vel"]
            for _ in range(3):
                C, _, h = prop.propagation(f=h, flags=flags, phi_obs=s["phi"], rho=C, vel=vel, is_convection=True, inflow_density=0)
                C = macro.macro_compute_C(dx=1.0, dt=1.0, h=h, C=C, flags=flags, vel=vel, mesh_grid=mesh)
                C = CLAMPS[spelling](C)
                h = col.collision(dx=1.0, dt=1.0, f=h, rho=C, vel=vel, flags=flags, force=None, mesh_grid=mesh,
                                  is_convection=True, KBC_type=int(api.KBCType.KBC_C))
            outs.append((h.cpu().numpy().copy(), np.asarray(C.cpu().numpy()).copy()))
        finally:
            L.ENABLED = True
    for a, b in zip(*outs):
        assert np.array_equal(a, b, equal_nan=True), spelling


WALL_ACTIONS = ["none"] * 6 + ["read_wall_f", "inplace_wall_f", "clone_wall_f", "read_vel_in", "skip_wall", "wall_twice", "read_rho",
                               "stale_stream_after"]


def _wall_program(api, seed, enabled, hw=None):
    """The four-call chain WITH solve_boundary (demo_2d_LBM_cylinder.py:138-172; stage A's axisymmetric form in a third of
    the programs), MRT / KBC / BGK collision objects, and user actions around the wall call."""
    rng = random.Random(1000 + seed)
    Hh, Ww = hw or (H, W)   # other extents reach the 1- and 4-cell-per-thread kernels (W odd / W % 4 == 0)
    L.ENABLED = enabled
    try:
        axisym = 1 if rng.random() < 0.35 else 0
        kind = "bgk" if axisym else rng.choice(["bgk", "mrt", "kbc"])
        P = api.SimulationParameters(dim=2, dtype=torch.float32, device=DEV, simulation_size=(1, 1, Hh, Ww), dt=1.0,
                                     contact_angle=None, Q=9, tau=0.8, axisymmetric_type=axisym, gravity_strength=1e-5)
        R = api.SimulationRunner(parameters=P)
        prop, macro, wall = R.create_propagation(), R.create_macro_compute(), R.create_LBM_fluid_solid_coupling()
        col = {"bgk": R.create_collision, "mrt": R.create_collision_MRT, "kbc": R.create_collision_KBC}[kind]()
        col.preset_KBC(dx=1.0, dt=1.0, tau=0.8)
        s = S.random_state((Hh, Ww), 20 + seed % 5, DEV, axisym)
        mesh = api.create_2d_meshgrid_tensor(size=(1, 1, Hh, Ww), device=DEV, dtype=torch.float32) + 0.5
        f, flags, rho, vel, phi, obs_vel = s["f"], s["flags"], s["rho"], s["vel"], s["phi"], s["obs_vel"]
        force = s["force"] if rng.random() < 0.5 else None
        kbc_type = int(rng.choice([api.KBCType.KBC_C, api.KBCType.KBC_D, api.KBCType.LBGK])) if kind == "kbc" else int(api.KBCType.KBC_C)
        seen, used = [], set()

        def obs(x):
            seen.append(np.asarray(x.detach().cpu().numpy()).copy())

        for step in range(rng.randint(2, 3)):
            rho_in, vel_in, f1 = prop.propagation(f=f, flags=flags, phi_obs=phi, rho=rho, vel=vel, inflow_density=1.0,
                                                  inflow_vel=torch.tensor([0.0, 0.02]))
            a = rng.choice(WALL_ACTIONS)
            used.add(a)
            if a == "read_vel_in":
                obs(vel_in[0, :, 2])
            if a == "skip_wall":
                f2 = f1
            else:
                f2 = wall.solve_boundary(dt=1.0, dx=1.0, rho=rho_in, vel=vel_in, flags=flags, f=f1, phi_obs=phi, obs_vel=obs_vel, tau=0.8)
                if a == "wall_twice":
                    f2 = wall.solve_boundary(dt=1.0, dx=1.0, rho=rho_in, vel=vel_in, flags=flags, f=f2, phi_obs=phi, obs_vel=obs_vel, tau=0.8)
            if a == "read_wall_f":
                obs(f2[0, 5])
            elif a == "inplace_wall_f":
                f2[:, 2, 1, 1] -= 0.002
            elif a == "clone_wall_f":
                f2 = f2.clone()
            rho, vel = macro.macro_compute(dx=1.0, dt=1.0, f=f2, rho=rho_in, vel=vel_in, flags=flags)
            if a == "read_rho":
                obs(rho.min())
            f = col.collision(dx=1.0, dt=1.0, f=f2, rho=rho, vel=vel, flags=flags, force=force,
                              mesh_grid=mesh if axisym else None, KBC_type=kbc_type)
            if a == "stale_stream_after":
                obs(f1[0, 1, 3])         # the pre-wall streamed populations, after the fused step has run
        return seen, [np.asarray(t.detach().cpu().numpy()).copy() for t in (f, rho, vel)], used
    finally:
        L.ENABLED = True


@pytest.mark.parametrize("block", range(3))
def test_random_wall_chain_programs_emu(emu_api, block):
    fused0 = L.stats["fused"]
    for seed in range(block * 15, block * 15 + 15):
        want_seen, want_final, used = _wall_program(emu_api, seed, enabled=False)
        got_seen, got_final, _ = _wall_program(emu_api, seed, enabled=True)
        assert len(want_seen) == len(got_seen), f"program {seed}"
        for i, (a, b) in enumerate(zip(want_seen, got_seen)):
            assert np.array_equal(a, b, equal_nan=True), f"wall program {seed}: observation {i} differs ({sorted(used)})"
        for i, (a, b) in enumerate(zip(want_final, got_final)):
            assert np.array_equal(a, b, equal_nan=True), f"wall program {seed}: final field {i} differs ({sorted(used)})"
    print(f"block {block}: 15 wall-chain programs, {L.stats['fused'] - fused0} steps fused")
    assert L.stats["fused"] - fused0 >= 5
