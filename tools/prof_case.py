"""Profiling driver (not part of the product): a few steps of ONE step variant on cuda:0, for `ncu`.

    python tools/prof_case.py CASE [STEPS] [GRID]
CASE: bgk | bgk_api (the reference's three calls, moments stored, zero force field as the demo passes) | mrt | kbc_a | kbc_d |
      c3 (cylinder, MRT + wall + contact ring) | c4 (scaled concentration step)"""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402

import bench  # noqa: E402
import workloads as WL  # noqa: E402
from src.LBM.simulation import SimulationParameters, SimulationRunner  # noqa: E402

case = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
G = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
dev = torch.device("cuda:0")
if case in ("bgk", "mrt", "kbc_a", "kbc_d", "bgk_api"):
    rho, vel = bench.shear_rows(G, torch.arange(G), dev)
    flags = torch.ones((1, 1, G, G), dtype=torch.uint8, device=dev)
    P = SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, G, G), dt=1.0, contact_angle=None, Q=9,
                             tau=bench.shear_tau(G))
    R = SimulationRunner(P)
    col = R.create_collision()
    f = col.get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
    if case == "bgk_api":
        prop, macro = R.create_propagation(), R.create_macro_compute()
        force = torch.zeros((1, 2, G, G), dtype=torch.float32, device=dev)
        for _ in range(steps):
            f2 = prop.propagation(f=f, flags=flags)
            rho, vel = macro.macro_compute(dx=1.0, dt=1.0, f=f2, rho=rho, vel=vel, flags=flags)
            f = col.collision(dx=1.0, dt=1.0, f=f2, rho=rho, vel=vel, flags=flags, force=force, KBC_type=None)
    else:
        op, kt = {"bgk": ("bgk", None), "mrt": ("mrt", None), "kbc_a": ("kbc", 133), "kbc_d": ("kbc", 138)}[case]
        fs = R.create_fused_step(operator=op, kbc_type=kt)
        f2 = torch.empty_like(f)
        for _ in range(steps):
            fs.step(f, flags, None, None, f_out=f2, store_moments=False)
            f, f2 = f2, f
elif case == "c3":
    cflags, phi, crho, cvel, inflow_vel = WL.cylinder_state(dev, G, G, 205, (2048.0, 2048.0))
    tau = WL.cylinder_tau(G, G)
    P = SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, G, G), dt=1.0, density_gas=0,
                             density_fluid=1.0, contact_angle=torch.Tensor([0.5 * math.pi]).to(dev), Q=9, tau=tau)
    R = SimulationRunner(P)
    col = R.create_collision_MRT()
    col.preset_KBC(dx=1.0, dt=1.0, tau=tau)
    fs = R.create_fused_step(operator="mrt")
    f0 = col.get_feq_(dx=1.0, dt=1.0, rho=crho, vel=cvel)
    b = [f0, torch.empty_like(f0), crho, torch.empty_like(crho), cvel, torch.empty_like(cvel)]
    zero = torch.zeros_like(cvel)
    for _ in range(steps):
        fs.step(b[0], cflags, b[2], b[4], f_out=b[1], rho_out=b[3], vel_out=b[5], inflow_density=1.0, inflow_vel=inflow_vel,
                use_wall=True, wall_phi=phi, obs_vel=zero)
        b[0], b[1], b[2], b[3], b[4], b[5] = b[1], b[0], b[3], b[2], b[5], b[4]
elif case == "c4":
    P = SimulationParameters(simulation_size=(1, 1, G, G), device=dev, contact_angle=None, axisymmetric_type=1)
    R = SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    flags, mesh, phi, vel, C = WL.c4_state(dev, G, G, torch.arange(G), G // 128)
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    st = [h, torch.empty_like(h), C, torch.empty_like(C)]
    for _ in range(steps):
        fs.step(st[0], flags, st[2], vel, f_out=st[1], rho_out=st[3], phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
        st[0], st[1], st[2], st[3] = st[1], st[0], st[3], st[2]
torch.cuda.synchronize()
print("done", case, steps, G)
