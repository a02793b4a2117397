"""Random four-call chains (tests/test_lazy_user_programs.py: BGK / MRT / KBC C, D, LBGK, with and without solve_boundary,
x-axis symmetric stage-A form, forces, inflow, user actions in between) on the PRODUCT -- deferred chain, host-emulated
kernels -- against the ORACLE's restatement of the reference behind the same solver-object API (oracle/oracle_api.py).
The lazy-layer tests compare the product with itself (deferred vs immediate); this one closes the loop to the reference's
arithmetic: <= 1e-5 relative L2 on every observation and final field (BASELINE north_star), bit-identical reported."""
import numpy as np
import pytest

from test_lazy_user_programs import _wall_program
from util import REL_L2_TOL, rel_l2


@pytest.mark.parametrize("block", range(3))
def test_random_wall_chains_match_oracle(emu_api, oracle_api, block):
    exact = total = 0
    for seed in range(block * 15, block * 15 + 15):
        hw = [(18, 14), (11, 17), (9, 32), (13, 24), (7, 21)][seed % 5]   # 2, 1, 4, 4, 1 cells per thread
        want_seen, want_final, used = _wall_program(oracle_api, seed, enabled=False, hw=hw)
        got_seen, got_final, _ = _wall_program(emu_api, seed, enabled=True, hw=hw)
        assert len(want_seen) == len(got_seen)
        for what, ws, gs in (("observation", want_seen, got_seen), ("final field", want_final, got_final)):
            for i, (w, g) in enumerate(zip(ws, gs)):
                w, g = np.asarray(w, np.float64), np.asarray(g, np.float64)
                assert np.array_equal(np.isfinite(w), np.isfinite(g)), f"program {seed}: {what} {i}: finite masks differ"
                m = np.isfinite(w)
                e = rel_l2(g[m], w[m]) if m.any() and np.abs(w[m]).max() > 0 else float(np.abs(g[m] - w[m]).max(initial=0.0))
                assert e <= REL_L2_TOL, f"program {seed}: {what} {i} differs from the oracle by {e:.3e} ({sorted(used)})"
                total += 1
                exact += int(np.array_equal(w, g, equal_nan=True))
    print(f"block {block}: {exact} of {total} arrays bit-identical to the oracle")


@pytest.mark.parametrize("block", range(2))
def test_random_plain_chains_match_oracle(emu_api, oracle_api, block):
    """The three-call chain programs (two interleaved solver sets, inflow, in-place user writes) against the oracle; programs
    that write into a recorded input raise in the deferred run by design and are skipped here."""
    from test_lazy_user_programs import _program
    exact = total = skipped = 0
    for seed in range(block * 20, block * 20 + 20):
        want_seen, want_final, used = _program(oracle_api, seed, enabled=False)
        try:
            got_seen, got_final, _ = _program(emu_api, seed, enabled=True)
        except RuntimeError as e:
            assert "modified in place" in str(e)
            skipped += 1
            continue
        assert len(want_seen) == len(got_seen)
        for what, ws, gs in (("observation", want_seen, got_seen), ("final field", want_final, got_final)):
            for i, (w, g) in enumerate(zip(ws, gs)):
                if np.asarray(w).dtype.kind in "iub":
                    assert np.array_equal(w, g), f"program {seed}: {what} {i} (integer) differs"
                    continue
                w, g = np.asarray(w, np.float64), np.asarray(g, np.float64)
                assert np.array_equal(np.isfinite(w), np.isfinite(g)), f"program {seed}: {what} {i}: finite masks differ"
                m = np.isfinite(w)
                e = rel_l2(g[m], w[m]) if m.any() and np.abs(w[m]).max() > 0 else float(np.abs(g[m] - w[m]).max(initial=0.0))
                assert e <= REL_L2_TOL, f"program {seed}: {what} {i} differs from the oracle by {e:.3e} ({sorted(used)})"
                total += 1
                exact += int(np.array_equal(w, g, equal_nan=True))
    print(f"block {block}: {exact} of {total} arrays bit-identical to the oracle, {skipped} programs skipped (documented error)")
    assert skipped <= 10


def _conc_program(api, seed, enabled):
    """The concentration chain (demo_2d_LBM_JRL_C_cylinder.py:313-340) with random extents, symmetry axis, batch, phi_obs,
    clamp and inflow density."""
    import random

    import torch

    import jrl_lazy as L
    import jrl_scenarios as S
    rng = random.Random(5000 + seed)
    L.ENABLED = enabled
    try:
        Hh, Ww = rng.choice([(18, 14), (11, 17), (9, 32), (13, 24), (16, 16)])
        axisym, B = rng.choice([0, 1, 1, 2]), rng.choice([1, 1, 2])
        dev = torch.device("cpu")
        P = api.SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(B, 1, Hh, Ww), dt=1.0,
                                     contact_angle=None, Q=9, tau=0.8, axisymmetric_type=axisym)
        R = api.SimulationRunner(parameters=P)
        prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
        col.preset_KBC(dx=1.0, dt=1.0, tau=0.8, tau_D=rng.choice([0.7, 0.9, 1.0]))
        s = S.random_state((Hh, Ww), 30 + seed % 4, dev, axisym, B=B)
        mesh = api.create_2d_meshgrid_tensor(size=(B, 1, Hh, Ww), device=dev, dtype=torch.float32) + 0.5
        h, C, flags, vel = s["f"], s["rho"] * 1.1, s["flags"], s["vel"]
        phi = s["phi"] if rng.random() < 0.6 else None
        clamp, dens = rng.random() < 0.6, rng.choice([0, 0.5])
        for _ in range(rng.randint(2, 4)):
            C, _, h = prop.propagation(f=h, flags=flags, phi_obs=phi, rho=C, vel=vel, is_convection=True, inflow_density=dens)
            C = macro.macro_compute_C(dx=1.0, dt=1.0, h=h, C=C, flags=flags, vel=vel, mesh_grid=mesh)
            if clamp:
                C = torch.clamp(C, 0, 1)
            h = col.collision(dx=1.0, dt=1.0, f=h, rho=C, vel=vel, flags=flags, force=None, mesh_grid=mesh,
                              is_convection=True, KBC_type=int(api.KBCType.KBC_C))
        return [np.asarray(t.detach().cpu().numpy()).copy() for t in (h, C)]
    finally:
        L.ENABLED = True


@pytest.mark.parametrize("block", range(2))
def test_random_concentration_chains_match_oracle(emu_api, oracle_api, block):
    import jrl_lazy as L
    exact = total = 0
    fused0 = L.stats["fused"]
    for seed in range(block * 15, block * 15 + 15):
        want = _conc_program(oracle_api, seed, enabled=False)
        got = _conc_program(emu_api, seed, enabled=True)
        for i, (w, g) in enumerate(zip(want, got)):
            w, g = np.asarray(w, np.float64), np.asarray(g, np.float64)
            assert np.array_equal(np.isfinite(w), np.isfinite(g)), f"program {seed}: field {i}: finite masks differ"
            m = np.isfinite(w)
            e = rel_l2(g[m], w[m])
            assert e <= REL_L2_TOL, f"concentration program {seed}: field {i} differs from the oracle by {e:.3e}"
            total += 1
            exact += int(np.array_equal(w, g, equal_nan=True))
    print(f"block {block}: {exact} of {total} arrays bit-identical to the oracle, {L.stats['fused'] - fused0} steps fused")
    assert L.stats["fused"] - fused0 >= 30
