"""Small cases for `compute-sanitizer --tool memcheck|racecheck` (VERDICT r1 missing #7): a few launches of
every kernel family on grids small enough for the tool -- the pipelined step (TMA ring + mbarriers) with mixed flags so
that tile mask, x-axis fast path and the work-list second launch all run; the register kernel with INFLOW cells (the
in-place rho / vel writes the reference races on) and OUTFLOW rows; the per-operator chain; the walk loops."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests")); sys.path.append(os.path.join(ROOT, "tools"))
import torch  # noqa: E402

import jrl_native  # noqa: E402
jrl_native.load()
import jrl_lazy  # noqa: E402
import jrl_scenarios as S  # noqa: E402
import workloads as WL  # noqa: E402

api, dev = S.api_from_src(), torch.device("cuda:0")
what = sys.argv[1] if len(sys.argv) > 1 else "all"
if what in ("all", "pipe"):
    # pipelined BGK with obstacles (W = 1024: two tiles per row), moments stored, 3 steps
    st = S.random_state((24, 1024), 5, dev)
    P = api.SimulationParameters(simulation_size=(1, 1, 24, 1024), device=dev, tau=0.8, contact_angle=None)
    fs = api.SimulationRunner(P).create_fused_step(operator="bgk")
    f, r, u = st["f"], st["rho"], st["vel"]
    for _ in range(3):
        f, r, u = fs.step(f, st["flags"], r, u, phi_obs=st["phi"], inflow_density=1.0)
    # pipelined concentration form on the scaled tip geometry (x-axis symmetry, work list), 1024 x 1024
    G = 1024
    flags, mesh, phi, vel, C = WL.c4_state(dev, G, G, torch.arange(G), G // 128)
    P = api.SimulationParameters(simulation_size=(1, 1, G, G), device=dev, contact_angle=None, axisymmetric_type=1)
    R = api.SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    for _ in range(2):
        h, C = fs.step(h, flags, C, vel, phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
    torch.cuda.synchronize()
    print("pipe ok", float(C.sum()))
if what in ("all", "chain"):
    jrl_lazy.ENABLED = False   # every reference call = its own launch: k_stream, k_wall, k_moments, k_collide, k_contact_ring
    S.cylinder(api, dev, res=(64, 48), D=8, steps=3, center=(24.0, 22.0))
    S.jrl_fluid(api, dev, res=(48, 32), steps=3, radius_obs=2.50000000000001, inflow_height=8.50000000000001, Rg=3, vel_obs=0.02,
                gravity_strength=2e-5)
    jrl_lazy.ENABLED = True    # fused behind the four calls + the walk loops below the ABI
    S.cylinder(api, dev, res=(64, 48), D=8, steps=3, center=(24.0, 22.0))
    S.jrl_fluid(api, dev, res=(48, 32), steps=3, radius_obs=2.50000000000001, inflow_height=8.50000000000001, Rg=3, vel_obs=0.02,
                gravity_strength=2e-5, device_tip="run")
    S.jrl_conc(api, dev, res=(48, 32), balance_steps=2, walk_steps=3, radius_obs=2.50000000000001, inflow_height=8.50000000000001,
               Rg=3, vel_obs_conv=0.03)
    torch.cuda.synchronize()
    print("chain ok")
