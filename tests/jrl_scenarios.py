"""Workload builders for the five BASELINE.json configs, written against the reference's
solver-object API (SimulationParameters / SimulationRunner / prop / macro / collision /
solid_fluid_coupling) and nothing else.

The same functions drive three back ends:
  * this repo's CUDA drop-in (`src.LBM` under jrl-lbm_b200/) -- tests -m gpu, bench.py;
  * the unmodified reference imported on the CPU (oracle/ref_cpu.py) -- oracle/gen_golden.py;
  * the numpy oracle behind the same class names (oracle/oracle_api.py) -- CPU tests.
so a parity test is literally "run the same script body twice".  Each builder mirrors the
body of a reference demo script (cited per function); file output, tqdm, PNG/MP4 writing and
the HDF5 hand-off are left out (off the per-step path).

An `api` is any object with attributes SimulationParameters, SimulationRunner, CellType,
KBCType, AxiSymmetricType, create_2d_meshgrid_tensor, UnionPhiObs.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch


def api_from_src():
    """Bundle whichever `src.LBM` is importable in this process (the drop-in or the reference)."""
    from src.LBM.simulation import SimulationParameters, SimulationRunner
    from src.LBM.utils import (AxiSymmetricType, CellType, KBCType, UnionPhiObs,
                               create_2d_meshgrid_tensor)
    return SimpleNamespace(SimulationParameters=SimulationParameters, SimulationRunner=SimulationRunner,
                           CellType=CellType, KBCType=KBCType, AxiSymmetricType=AxiSymmetricType,
                           create_2d_meshgrid_tensor=create_2d_meshgrid_tensor, UnionPhiObs=UnionPhiObs)


def _np(t):
    return t.detach().cpu().numpy().copy()


# ------------------------------------------------------------------------------------------------
# C1 / C2: periodic double shear layer   (demo/demo_2d_LBM.py:17-119, demo_2d_LBM_KBC.py:17-120)
# ------------------------------------------------------------------------------------------------

def shear_layer_state(res, device, dtype=torch.float32, rho0=0.265, Vmax=0.2, sigma=0.05, kappa=80.0):
    """Analytic initial rho/u of demo_2d_LBM.py:84-98, vectorised (the demo fills it with a
    Python double loop in float64 `math` calls and stores to fp32; so does this)."""
    H, W = res
    j = torch.arange(H, dtype=torch.float64).view(H, 1)
    i = torch.arange(W, dtype=torch.float64).view(1, W)
    uy = (sigma * Vmax * torch.sin(2.0 * math.pi * (1.0 * i / W + 0.25))).expand(H, W)
    ux = torch.where(j <= (H / 2.0), Vmax * torch.tanh(kappa * (1.0 * j / H - 0.25)),
                     Vmax * torch.tanh(kappa * (-1.0 * j / H + 0.75))).expand(H, W)
    vel = torch.stack([ux, uy]).unsqueeze(0).to(dtype).to(device).contiguous()
    rho = torch.full((1, 1, H, W), rho0, dtype=dtype, device=device)
    return rho, vel


def shear_layer(api, device, res=(130, 130), steps=1000, operator="bgk", kbc_type=None,
                Re=3000.0, dx=1.0, dt=1.0, snapshots=(), fused=False):
    """operator: 'bgk' (demo_2d_LBM.py), 'mrt' (what demo_2d_LBM_KBC.py really runs)
    or 'kbc' (LBMCollisionKBC2d with `kbc_type`).  Returns final fields as numpy arrays, plus
    `snap_<n>_*` copies after each step count listed in `snapshots`."""
    dtype = torch.float32
    c = dx / dt
    cs2 = c * c / 3.0
    Vmax = 0.2
    tau = 0.5 + (Vmax * max(res) * dx / Re) / cs2
    P = api.SimulationParameters(
        dim=2, dtype=dtype, device=device, simulation_size=(1, 1, *res), dt=dt, density_gas=0.038,
        density_fluid=0.265, contact_angle=torch.Tensor([0.5 * math.pi]).to(device).to(dtype), Q=9,
        tau=tau, k=0.33)
    R = api.SimulationRunner(parameters=P)
    prop = R.create_propagation()
    macro = R.create_macro_compute()
    if operator == "bgk":
        collision = R.create_collision()
    elif operator == "mrt":
        collision = R.create_collision_MRT()
        collision.preset_KBC(dx=dx, dt=dt, tau=tau)
    elif operator == "kbc":
        collision = R.create_collision_KBC()
        collision.preset_KBC(dx=dx, dt=dt, tau=tau)
    else:
        raise ValueError(operator)
    flags = torch.full((1, 1, *res), int(api.CellType.FLUID), dtype=torch.uint8, device=device)
    density, vel = shear_layer_state(res, device, dtype)
    force = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    f = collision.get_feq_(dx=dx, dt=dt, rho=density, vel=vel, force=force)
    out = {"tau": tau, "f0": _np(f)}
    if fused:  # extension of this repo: the same loop body as ONE launch per step (jrl_step.FusedStep)
        fs = R.create_fused_step(operator=operator, kbc_type=kbc_type if operator == "kbc" else None, dx=dx, dt=dt)
    if fused == "run":  # the whole loop below the C ABI (jrl_run_f32), cut only where a snapshot is wanted
        done = 0
        for stop in sorted(set(int(x) for x in snapshots if 0 < int(x) < steps)) + [steps]:
            f, density, vel = fs.run(f, flags, density, vel, stop - done, force=force)
            done = stop
            if done in snapshots:
                out[f"snap_{done}_rho"] = _np(density)
                out[f"snap_{done}_vel"] = _np(vel)
        steps = 0
    for step in range(steps):
        if fused:
            f, density, vel = fs.step(f, flags, density, vel, force=force)
        else:
            f = prop.propagation(f=f, flags=flags)
            density, vel = macro.macro_compute(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags)
            f = collision.collision(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags, force=force,
                                    KBC_type=kbc_type)
        R.step()
        if (step + 1) in snapshots:
            out[f"snap_{step + 1}_rho"] = _np(density)
            out[f"snap_{step + 1}_vel"] = _np(vel)
    out.update(rho=_np(density), vel=_np(vel), f=_np(f), vort=_np(macro.get_vort(vel=vel, dx=dx)))
    return out


# ------------------------------------------------------------------------------------------------
# C3: flow past a cylinder   (demo/demo_2d_LBM_cylinder.py:26-173)
# ------------------------------------------------------------------------------------------------

def cylinder_phi(mesh_grid, center, radius):
    """Signed distance to the cylinder surface at cell centres, demo_2d_LBM_cylinder.py:26-40."""
    x = mesh_grid[:, 0:1]
    y = mesh_grid[:, 1:2]
    d2 = (x - center[0]) * (x - center[0]) + (y - center[1]) * (y - center[1]) - radius * radius
    return torch.where(d2 > 0, torch.sqrt(d2), -torch.sqrt(-d2))


def cylinder(api, device, res=(800, 400), D=20, steps=2000, Re=5000.0, Vmax=0.05, dx=1.0, dt=1.0,
             center=None, snapshots=(), fused=False):
    """res = [H, W]; inflow along +y from row 0, outflow at row H-1, OBSTACLE ring + cylinder,
    MRT collision with the KBC_C flag ignored, moving-wall coupling with a resting wall."""
    dtype = torch.float32
    c = dx / dt
    cs2 = c * c / 3.0
    tau = 0.5 + (Vmax * max(res) * dx / Re) / cs2
    P = api.SimulationParameters(
        dim=2, dtype=dtype, device=device, simulation_size=(1, 1, *res), dt=dt, density_gas=0,
        density_fluid=1.0, contact_angle=torch.Tensor([0.5 * math.pi]).to(device).to(dtype), Q=9,
        tau=tau, k=0.33)
    R = api.SimulationRunner(parameters=P)
    prop = R.create_propagation()
    macro = R.create_macro_compute()
    collision = R.create_collision_MRT()
    collision.preset_KBC(dx=dx, dt=dt, tau=tau)
    coupling = R.create_LBM_fluid_solid_coupling()
    CT = api.CellType
    flags = torch.full((1, 1, *res), int(CT.OBSTACLE), dtype=torch.uint8, device=device)
    flags[..., 1:-1, 1:-1] = int(CT.FLUID)
    flags[..., 0, 1:-1] = int(CT.INFLOW)
    flags[..., -1, 1:-1] = int(CT.OUTFLOW)
    inflow_vel = torch.Tensor([0.0, Vmax]).to(device).to(dtype)
    density = torch.full((1, 1, *res), 1.0, dtype=dtype, device=device)
    vel = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    force = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    phi_obs_vel = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    if center is None:
        center = (10.0 * D, 10.0 * D)
    mesh_grid = api.create_2d_meshgrid_tensor(size=(1, 1, *res), device=device, dtype=dtype) + 0.5 * dx
    phi_obs = cylinder_phi(mesh_grid, center, D / 2.0)
    flags[phi_obs <= 0] = int(CT.OBSTACLE)
    vel = torch.where(flags == int(CT.INFLOW), inflow_vel.reshape(-1, 2, 1, 1), vel)
    f = collision.get_feq_(dx=dx, dt=dt, rho=density, vel=vel, force=force)
    out = {"tau": tau, "flags": _np(flags), "phi_obs": _np(phi_obs)}
    if fused:
        fs = R.create_fused_step(operator="mrt", dx=dx, dt=dt)
    if fused == "run":
        done = 0
        for stop in sorted(set(int(x) for x in snapshots if 0 < int(x) < steps)) + [steps]:
            f, density, vel = fs.run(f, flags, density, vel, stop - done, force=force, inflow_density=1.0,
                                     inflow_vel=inflow_vel, use_wall=True, wall_phi=phi_obs, obs_vel=phi_obs_vel)
            done = stop
            if done in snapshots:
                out[f"snap_{done}_rho"] = _np(density)
                out[f"snap_{done}_vel"] = _np(vel)
        steps = 0
    for step in range(steps):
        if fused:
            f, density, vel = fs.step(f, flags, density, vel, force=force, inflow_density=1.0, inflow_vel=inflow_vel,
                                      use_wall=True, wall_phi=phi_obs, obs_vel=phi_obs_vel)
        else:
            density, vel, f = prop.propagation(f=f, flags=flags, rho=density, vel=vel, inflow_density=1.0,
                                               inflow_vel=inflow_vel)
            f = coupling.solve_boundary(dt=dt, dx=dx, rho=density, vel=vel, flags=flags, f=f, phi_obs=phi_obs,
                                        obs_vel=phi_obs_vel, tau=tau)
            density, vel = macro.macro_compute(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags)
            f = collision.collision(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags, force=force,
                                    KBC_type=int(api.KBCType.KBC_C))
        R.step()
        if (step + 1) in snapshots:
            out[f"snap_{step + 1}_rho"] = _np(density)
            out[f"snap_{step + 1}_vel"] = _np(vel)
    out.update(rho=_np(density), vel=_np(vel), f=_np(f))
    return out


# ------------------------------------------------------------------------------------------------
# C4 / C5 stage A: axisymmetric flow around an approaching probe tip
#   (demo/demo_2d_LBM_JRL_fluid_cylinder.py:39-268)
# ------------------------------------------------------------------------------------------------

def tip_phi(api, mesh_grid, box_center, box_radius, is_negative=True, geom="cylinder"):
    """CalculateObsPhi / CalculateObsPhi_forC: box signed distance (max norm) united with the
    top wall, demo_2d_LBM_JRL_fluid_cylinder.py:39-49, demo_2d_LBM_JRL_C_cylinder.py:31-41 (the infinite-plane demos
    use the same function with Rg = 200); geom="45degree": the cone of demo_2d_LBM_JRL_*_45degree.py:31-46."""
    res_y = mesh_grid.shape[-2]
    if geom == "45degree":
        x, y = mesh_grid[:, 0:1], mesh_grid[:, 1:2]
        phi = (y - box_center[1]) - box_radius
        x1 = (box_center + box_radius)[0]
        y1 = (box_center + box_radius)[1]
        A_, B_, C_ = 1.0, 1.0, -(x1 + y1)
        phi = torch.maximum(phi, (A_ * x + B_ * y + C_) / math.sqrt(2))
        if is_negative:
            phi = api.UnionPhiObs(phi, -(y - res_y), alpha=1)
        return phi
    tx = (mesh_grid[:, 0:1] - box_center[0]) - box_radius
    ty = (mesh_grid[:, 1:2] - box_center[1]) - box_radius
    phi = torch.maximum(tx, ty)
    if is_negative:
        phi = api.UnionPhiObs(phi, -(mesh_grid[:, 1:2] - res_y), alpha=1)
    return phi


def jrl_fluid(api, device, res=(128, 64), steps=200, Re=3e-4, gravity_strength=0.0, radius_obs=6.50000000000001,
              inflow_height=20.50000000000001, Rg=4, vel_obs=None, x_interval=0.1, dx=1, dt=1, tau=1.0,
              snapshots=(), device_tip=False, geom="cylinder"):
    """Stage A with the demo's per-step host work (phi / flags / wall velocity restamp) kept as
    torch ops.  `vel_obs` defaults to Re*visc/radius_obs as in the demo (:78); tests pass a
    larger value so the tip crosses cells within a few hundred steps.  `gravity_strength` is
    the lattice-unit value (:91)."""
    dtype = torch.float32
    c = dx / dt
    cs2 = c * c / 3.0
    visc = cs2 * (tau - 0.5)
    if vel_obs is None:
        vel_obs = Re * visc / radius_obs
    inflow_height_int = int(inflow_height + 0.5 * dx)
    P = api.SimulationParameters(
        dim=2, dtype=dtype, device=device, simulation_size=(1, 1, *res), dt=dt, density_fluid=1,
        axisymmetric_type=int(api.AxiSymmetricType.LINE_X_EQ_0), contact_angle=None, Q=9, tau=tau,
        gravity_strength=gravity_strength)
    R = api.SimulationRunner(parameters=P)
    CT = api.CellType
    flags = torch.zeros((1, 1, *res), dtype=torch.uint8, device=device)
    vel = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    density = torch.zeros((1, 1, *res), dtype=dtype, device=device)
    mesh_grid = api.create_2d_meshgrid_tensor(size=(1, 1, *res), device=device, dtype=dtype) + 0.5 * dx
    prop = R.create_propagation()
    macro = R.create_macro_compute()
    collision_f = R.create_collision()
    collision_f.preset_KBC(dx=dx, dt=dt, tau=tau, tau_D=None)
    coupling = R.create_LBM_fluid_solid_coupling()
    Rg_radius = int(Rg * radius_obs + 0.5 * dx)
    flags[...] = int(CT.OUTFLOW)
    flags[..., 1:-1, 1:-1] = int(CT.FLUID)
    flags[..., 0] = int(CT.FLUID)
    flags[..., -1, :] = int(CT.OBSTACLE)
    flags[..., 0:inflow_height_int, 0:Rg_radius] = int(CT.OBSTACLE)
    density[flags == int(CT.OBSTACLE)] = 1
    density[flags != int(CT.OBSTACLE)] = 1
    force = density * collision_f._gravity
    f = collision_f.get_feq_(dx=dx, dt=dt, rho=density, vel=vel, tau=tau, force=force)
    box_radius = Rg_radius + 100.0
    walks = inflow_height
    walks_int = inflow_height_int
    box_center = torch.Tensor([Rg_radius - box_radius, walks - box_radius]).to(device).to(dtype)
    if geom != "45degree":  # demo_2d_LBM_JRL_fluid_cylinder.py:193-195; the 45-degree demo stamps in the loop only
        flags[..., 0:walks_int, 0:Rg_radius] = int(CT.OBSTACLE)
        vel[..., 0, 0:walks_int, 0:Rg_radius] = 0
        vel[..., 1, 0:walks_int, 0:Rg_radius] = vel_obs
    out = {}
    if device_tip == "run":  # extension of this repo: the whole loop below the C ABI (jrl_fluid_walk_f32)
        import jrl_step
        done = 0
        for stop in sorted(set(int(x) for x in snapshots if 0 < int(x) < steps)) + [steps]:
            f, density, vel, walks, phi_obs, phi_obs_vel = jrl_step.fluid_walk(
                collision_f, f, flags, density, vel, mesh_grid, force, walks, stop - done, box_center_x=float(box_center[0]),
                box_radius=box_radius, Rg_radius=Rg_radius, vel_obs=vel_obs, tau=tau, dx=dx, dt=dt,
                shape=1 if geom == "45degree" else 0)
            done = stop
            if done in snapshots:
                out[f"snap_{done}_rho"] = _np(density)
                out[f"snap_{done}_vel"] = _np(vel)
        steps = 0
    for step in range(steps):
        walks_int = int(walks + 0.5 * dx)
        box_center[1] = walks - box_radius
        phi_obs = tip_phi(api, mesh_grid, box_center, box_radius, geom=geom)
        if geom == "45degree":                                            # demo_2d_LBM_JRL_fluid_45degree.py:203-220
            is_obs = phi_obs <= 0
            phi_obs_vel = torch.where(is_obs.repeat(1, 2, 1, 1),
                                      torch.Tensor([0, vel_obs]).to(device).to(dtype).reshape(1, 2, 1, 1).expand_as(vel),
                                      torch.zeros_like(vel))
            phi_obs_vel[:, :, -1, :] = 0
            flags[is_obs] = int(CT.OBSTACLE)
            vel = torch.where(is_obs.repeat(1, 2, 1, 1), phi_obs_vel, vel)
        else:
            phi_obs_vel = torch.zeros_like(vel)
            phi_obs_vel[..., 1, 0:walks_int, 0:Rg_radius] = vel_obs
            flags[..., 0:walks_int, 0:Rg_radius] = int(CT.OBSTACLE)
            vel[..., 0, 0:walks_int, 0:Rg_radius] = 0
            vel[..., 1, 0:walks_int, 0:Rg_radius] = vel_obs
        f = prop.propagation(f=f, flags=flags, phi_obs=phi_obs, rho=density, vel=vel)
        f = coupling.solve_boundary(dt=dt, dx=dx, rho=density, vel=vel, flags=flags, f=f, phi_obs=phi_obs,
                                    obs_vel=phi_obs_vel, tau=tau)
        density, vel = macro.macro_compute(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags)
        f = collision_f.collision(dx=dx, dt=dt, f=f, rho=density, vel=vel, flags=flags, mesh_grid=mesh_grid,
                                  force=force)
        walks += vel_obs * dt
        if (step + 1) in snapshots:
            out[f"snap_{step + 1}_rho"] = _np(density)
            out[f"snap_{step + 1}_vel"] = _np(vel)
    out.update(rho=_np(density), vel=_np(vel), f=_np(f), flags=_np(flags), phi_obs=_np(phi_obs), walks=walks)
    return out


# ------------------------------------------------------------------------------------------------
# C4 / C5 stage B: concentration field + tip-current records
#   (demo/demo_2d_LBM_JRL_C_cylinder.py:44-407)
# ------------------------------------------------------------------------------------------------

def analytic_tip_velocity(mesh_grid, walks, r0):
    """Stand-in for the stage-A velocity field (SURVEY.md section 8d, C4 option (b)): a smooth,
    divergence-free-ish decaying flow around the tip, normalised to the tip speed.
    u = (-x/(2 r0), 1) * exp(-d/r0), d = distance of the cell centre from the tip face centre."""
    x = mesh_grid[:, 0:1]
    y = mesh_grid[:, 1:2]
    d = torch.sqrt(x * x + (y - walks) * (y - walks))
    w = torch.exp(-d / r0)
    return torch.cat([-x / (2.0 * r0) * w, w], dim=1)


def jrl_conc(api, device, res=(128, 64), balance_steps=200, walk_steps=200, Pe=0.34013, radius_obs=6.50000000000001,
             inflow_height=20.50000000000001, Rg=4, vel_obs_conv=None, is_negative=True, clamp=True,
             dx=1, dt=1, vel_field=None, fused=False, vel_snaps=None, x_interval=0.1, device_tip=False, geom="cylinder"):
    """Stage B.  `vel_field`: callable(mesh_grid, walks) -> normalised [1,2,H,W] velocity
    (defaults to `analytic_tip_velocity`); the demo interpolates HDF5 snapshots instead
    (:294-303).  `vel_snaps` [n,2,H,W]: velocity snapshots at tip travel 0, x_interval, 2 x_interval, ...
    interpolated per step exactly as the demo does (:294-303); `device_tip`: the walk loop's host work
    (phi, flags, C reset, C_vel) as ONE launch (jrl_solver.tip_prepare_conc) instead of torch ops; needs
    vel_snaps.  Returns final C, h and the per-step inflow/current flux series (:344-362)."""
    dtype = torch.float32
    c = dx / dt
    cs2 = c * c / 3.0
    inflow_height_int = int(inflow_height + 0.5 * dx)
    tau_D = 1.0
    Dc = cs2 * (tau_D - 0.5)
    if vel_obs_conv is None:
        vel_obs_conv = Pe * Dc / radius_obs
    if vel_obs_conv > 0.1:
        vel_obs_conv = 0.1
        Dc = vel_obs_conv * radius_obs / Pe
        tau_D = 0.5 + Dc / cs2
    P = api.SimulationParameters(
        dim=2, dtype=dtype, device=device, simulation_size=(1, 1, *res), dt=dt,
        axisymmetric_type=int(api.AxiSymmetricType.LINE_X_EQ_0), contact_angle=None, Q=9)
    R = api.SimulationRunner(parameters=P)
    CT = api.CellType
    C_flags = torch.zeros((1, 1, *res), dtype=torch.uint8, device=device)
    C = torch.ones((1, 1, *res), dtype=dtype, device=device)
    C_vel = torch.zeros((1, 2, *res), dtype=dtype, device=device)
    mesh_grid = api.create_2d_meshgrid_tensor(size=(1, 1, *res), device=device, dtype=dtype) + 0.5 * dx
    prop = R.create_propagation()
    macro = R.create_macro_compute()
    collision_h = R.create_collision()
    collision_h.preset_KBC(dx=dx, dt=dt, tau=1.0, tau_D=tau_D)
    Rg_radius = int(Rg * radius_obs + 0.5 * dx)
    radius_obs_int = int(radius_obs + 0.5 * dx)
    C_flags[...] = int(CT.OUTFLOW)
    C_flags[..., 1:-1, 1:-1] = int(CT.FLUID)
    C_flags[..., 0] = int(CT.FLUID)
    C_flags[..., -1, :] = int(CT.OBSTACLE) if is_negative else int(CT.INFLOW_2)
    C_flags[..., :-1, -1] = int(CT.INFLOW_2) if (is_negative or geom != "45degree") else int(CT.OUTFLOW)
    C_flags[..., 0, Rg_radius:] = int(CT.INFLOW_2)
    C_flags[..., 0:inflow_height_int, 0:Rg_radius] = int(CT.OBSTACLE)
    C_flags[..., inflow_height_int - 1, 0:radius_obs_int] = int(CT.INFLOW)
    area = math.pi * torch.Tensor([i * i for i in range(radius_obs_int + 1)]).to(device).to(dtype)
    area = area[1:] - area[:-1]
    h = collision_h.get_feq_(dx=dx, dt=dt, rho=C, vel=C_vel, tau=tau_D)
    box_radius = Rg_radius + 100.0
    walks = inflow_height
    walks_int = inflow_height_int
    box_center = torch.Tensor([Rg_radius - box_radius, walks - box_radius]).to(device).to(dtype)
    if geom == "45degree":                                                # demo_2d_LBM_JRL_C_45degree.py:211-216
        C_flags[tip_phi(api, mesh_grid, box_center, box_radius, is_negative, geom) <= 0] = int(CT.OBSTACLE)
    else:
        C_flags[..., 0:walks_int, 0:Rg_radius] = int(CT.OBSTACLE)
    C_flags[..., walks_int - 1, 0:radius_obs_int] = int(CT.INFLOW)
    C[C_flags == int(CT.OBSTACLE)] = 0.0
    balance = []
    if fused:
        fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=tau_D, dx=dx, dt=dt)
    for step in range(balance_steps):                                     # :215-241
        C_vel[...] = 0
        if fused:
            h, C = fs.step(h, C_flags, C, C_vel, mesh_grid=mesh_grid, inflow_density=0)
        else:
            C, _, h = prop.propagation(f=h, flags=C_flags, rho=C, vel=C_vel, is_convection=True, inflow_density=0)
            C = macro.macro_compute_C(dx=dx, dt=dt, h=h, C=C, flags=C_flags, vel=C_vel, mesh_grid=mesh_grid)
            h = collision_h.collision(dx=dx, dt=dt, f=h, rho=C, vel=C_vel, flags=C_flags, force=None,
                                      mesh_grid=mesh_grid, is_convection=True, KBC_type=int(api.KBCType.KBC_C))
        balance.append(((C)[..., walks_int, 0:radius_obs_int].flatten() * area).sum())
    out = {"C_balance": _np(C), "h_balance": _np(h), "tau_D": tau_D, "vel_obs_conv": vel_obs_conv}
    if vel_field is None:
        vel_field = lambda mg, w: analytic_tip_velocity(mg, w, radius_obs)  # noqa: E731
    inflow, current = [], []
    if device_tip == "run":  # the whole walk loop below the C ABI (jrl_conc_walk_f32)
        import jrl_step
        h, C, rec, walks, phi_obs, C_vel = jrl_step.conc_walk(
            fs, h, C_flags, C, mesh_grid, vel_snaps, walks, walk_steps, box_center_x=float(box_center[0]), box_radius=box_radius,
            Rg_radius=Rg_radius, radius_obs_int=radius_obs_int, vel_obs_conv=vel_obs_conv, inflow_height=inflow_height,
            x_interval=x_interval, is_negative=is_negative, clamp01=clamp, dx=dx, dt=dt,
            shape=1 if geom == "45degree" else 0)
        col = 1 if geom == "45degree" else 0   # the 45-degree demo records the plain strip sums (:353-376)
        inflow, current = list(rec[:, col]), list(rec[:, 2 + col])
        for _ in range(walk_steps):
            R.step()
        walk_steps = 0
    if device_tip:
        import jrl_solver
        phi_buf, vel_buf = torch.empty_like(C), torch.empty_like(C_vel)
    for step in range(walk_steps):                                        # :281-407
        walks_int = int(walks + 0.5 * dx)
        box_center[1] = walks - box_radius
        if vel_snaps is not None:                                         # :294-303
            x_comp = walks - inflow_height
            batch_idx = min(int(x_comp / x_interval), vel_snaps.shape[0] - 1)
            t_w = x_comp / x_interval - batch_idx if batch_idx + 1 < vel_snaps.shape[0] else None
        if device_tip:
            phi_obs, C_vel = jrl_solver.tip_prepare_conc(
                mesh_grid, C_flags, C, vel_snaps[batch_idx], None if t_w is None else vel_snaps[batch_idx + 1],
                0.0 if t_w is None else t_w, box_center, box_radius, walks_int, Rg_radius, radius_obs_int, vel_obs_conv,
                is_negative=is_negative, phi_out=phi_buf, vel_out=vel_buf, shape=1 if geom == "45degree" else 0)
        else:
            phi_obs = tip_phi(api, mesh_grid, box_center, box_radius, is_negative, geom)
            if not is_negative:
                C_flags[..., -1, :] = int(CT.INFLOW_2)
            C_flags[phi_obs < 0] = int(CT.OBSTACLE)
            C_flags[..., walks_int - 1, 0:radius_obs_int] = int(CT.INFLOW)
            C[C_flags == int(CT.OBSTACLE)] = 0.0
            if vel_snaps is not None:
                if t_w is not None:
                    vel_now = vel_snaps[batch_idx] * (1 - t_w) + vel_snaps[batch_idx + 1] * t_w
                else:
                    vel_now = vel_snaps[batch_idx]
                C_vel = vel_now.unsqueeze(0) / 1 * vel_obs_conv
            else:
                C_vel = vel_field(mesh_grid, walks) * vel_obs_conv
            C_vel[:, 0, 0:walks_int, 0:Rg_radius] = 0
            C_vel[:, 1, 0:walks_int, 0:Rg_radius] = vel_obs_conv
        if fused:
            h, C = fs.step(h, C_flags, C, C_vel.contiguous(), phi_obs=phi_obs, mesh_grid=mesh_grid, inflow_density=0,
                           clamp01=clamp)
        else:
            C, _, h = prop.propagation(f=h, flags=C_flags, phi_obs=phi_obs, rho=C, vel=C_vel, is_convection=True,
                                       inflow_density=0)
            C = macro.macro_compute_C(dx=dx, dt=dt, h=h, C=C, flags=C_flags, vel=C_vel, mesh_grid=mesh_grid)
            if clamp:
                C = torch.clamp(C, 0, 1)
            h = collision_h.collision(dx=dx, dt=dt, f=h, rho=C, vel=C_vel, flags=C_flags, force=None,
                                      mesh_grid=mesh_grid, is_convection=True, KBC_type=int(api.KBCType.KBC_C))
        for y0, series in ((walks_int, inflow), (walks_int + 1, current)):
            if geom == "45degree":                                        # demo_2d_LBM_JRL_C_45degree.py:353-376
                series.append(((C)[..., y0, 0:radius_obs_int].flatten() * area).sum())
                continue
            dC_z = ((C)[..., y0, 0:radius_obs_int] - (C)[..., y0 - 1, 0:radius_obs_int]).flatten()
            r = mesh_grid[..., 0:1, y0, 0:radius_obs_int + 1].flatten()
            rC = r * (C)[..., y0, 0:radius_obs_int + 1].flatten()
            dC_r = (rC[1:] - rC[:-1]) / r[:-1]
            series.append(((dC_z + dC_r) * area).sum())
        R.step()
        walks += vel_obs_conv * dt
    out.update(C=_np(C), h=_np(h), flags=_np(C_flags),
               balance=torch.stack(balance).cpu().numpy() if balance else None,
               inflow=torch.stack(inflow).cpu().numpy() if inflow else None,
               current=torch.stack(current).cpu().numpy() if current else None, walks=walks)
    return out


# ------------------------------------------------------------------------------------------------
# Single-call coverage of every API method on seeded random inputs (no reference demo does
# this; it is the per-operator parity harness of SURVEY.md section 8c).
# ------------------------------------------------------------------------------------------------

def random_state(res, seed, device, axisym=0, obstacles=True, io_cells=True, B=1):
    """Seeded random flags/f/phi/rho/vel/force with every CellType present."""
    g = torch.Generator().manual_seed(seed)
    H, W = res
    r = torch.rand((B, 1, H, W), generator=g)
    flags = torch.full((B, 1, H, W), 1, dtype=torch.uint8)
    if obstacles:
        flags[r < 0.15] = 2
        flags[(r >= 0.15) & (r < 0.18)] = 0
        flags[(r >= 0.18) & (r < 0.20)] = 4
    if io_cells:
        flags[(r >= 0.20) & (r < 0.23)] = 8
        flags[(r >= 0.23) & (r < 0.26)] = 32
        flags[:, :, 0, 2:-2:3] = 16
        flags[:, :, 2:-2:2, W - 1] = 16
    f = torch.rand((B, 9, H, W), generator=g) * 0.2 + 0.05
    phi = torch.rand((B, 1, H, W), generator=g) * 1.6 - 0.3
    rho = torch.rand((B, 1, H, W), generator=g) * 0.5 + 0.75
    vel = (torch.rand((B, 2, H, W), generator=g) - 0.5) * 0.2
    force = (torch.rand((B, 2, H, W), generator=g) - 0.5) * 0.01
    obs_vel = (torch.rand((B, 2, H, W), generator=g) - 0.5) * 0.05
    return {k: v.to(device).contiguous() for k, v in
            dict(flags=flags, f=f, phi=phi, rho=rho, vel=vel, force=force, obs_vel=obs_vel).items()}


def single_calls(api, device, res=(20, 24), seed=0, axisym=0, tau=0.8, tau_D=0.9, contact_angle=None,
                 dx=1.0, dt=1.0, B=1):
    """Call every solver-object method once on the same random state; return all outputs."""
    dtype = torch.float32
    s = random_state(res, seed, device, axisym, B=B)
    ca = None if contact_angle is None else torch.Tensor([contact_angle]).to(device).to(dtype)
    P = api.SimulationParameters(dim=2, dtype=dtype, device=device, simulation_size=(B, 1, *res), dt=dt,
                                 axisymmetric_type=axisym, contact_angle=ca, Q=9, tau=tau,
                                 gravity_strength=1e-4)
    R = api.SimulationRunner(parameters=P)
    prop, macro, coupling = R.create_propagation(), R.create_macro_compute(), R.create_LBM_fluid_solid_coupling()
    bgk, mrt, kbc = R.create_collision(), R.create_collision_MRT(), R.create_collision_KBC()
    for c in (bgk, mrt, kbc):
        c.preset_KBC(dx=dx, dt=dt, tau=tau, tau_D=tau_D)
    mesh = api.create_2d_meshgrid_tensor(size=(B, 1, *res), device=device, dtype=dtype) + 0.5 * dx
    flags, f, phi, rho, vel, force, obs_vel = (s[k] for k in ("flags", "f", "phi", "rho", "vel", "force", "obs_vel"))
    out = {}
    # streaming, all argument combinations the demos use
    out["stream_plain"] = _np(prop.propagation(f=f, flags=flags) if not bool((flags == 32).any())
                              else prop.propagation(f=f, flags=flags)[2])
    r1, v1 = rho.clone(), vel.clone()
    res3 = prop.propagation(f=f, flags=flags, phi_obs=phi, rho=r1, vel=v1)
    out["stream_phi"] = _np(res3[2] if isinstance(res3, tuple) else res3)
    out["stream_phi_rho"], out["stream_phi_vel"] = _np(r1), _np(v1)
    r2, v2 = rho.clone(), vel.clone()
    inflow_vel = torch.Tensor([0.01, 0.05]).to(device).to(dtype)
    rr, vv, ff = prop.propagation(f=f, flags=flags, rho=r2, vel=v2, inflow_density=1.0, inflow_vel=inflow_vel)
    out["stream_inflow"], out["stream_inflow_rho"], out["stream_inflow_vel"] = _np(ff), _np(rr), _np(vv)
    r3, v3 = rho.clone(), vel.clone()
    rr, vv, ff = prop.propagation(f=f, flags=flags, phi_obs=phi, rho=r3, vel=v3, is_convection=True, inflow_density=0)
    out["stream_conc"], out["stream_conc_rho"] = _np(ff), _np(rr)
    out["rebounce"] = _np(prop.rebounce_obstacle(f, flags))
    # moving wall
    out["wall_moving"] = _np(coupling.solve_boundary(dt=dt, dx=dx, rho=rho, vel=vel, flags=flags, f=f, phi_obs=phi,
                                                    obs_vel=obs_vel, tau=tau))
    out["wall_rest"] = _np(coupling.solve_boundary(dt=dt, dx=dx, rho=rho, vel=vel, flags=flags, f=f, phi_obs=phi,
                                                  obs_vel=None))
    # moments
    rn, vn = macro.macro_compute(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags)
    out["macro_rho"], out["macro_vel"] = _np(rn), _np(vn)
    rk, vk = macro.ApplyCompute_fluid(dx, dt, f, rho, vel, flags)
    out["macroK_rho"], out["macroK_vel"] = _np(rk), _np(vk)
    out["macro_C"] = _np(macro.macro_compute_C(dx=dx, dt=dt, h=f, C=rho, flags=flags, vel=vel, mesh_grid=mesh))
    out["macroK_C"] = _np(macro.ApplyCompute_C(dx, dt, f, rho, flags, vel, mesh))
    out["vort"] = _np(macro.get_vort(vel=vel, dx=dx))
    # equilibrium, stencils
    out["feq"] = _np(bgk.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel))
    out["feq_force"] = _np(bgk.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel, force=force))
    out["grad"] = _np(bgk.get_grad(rho, dx=dx, flags=flags))
    out["div"] = _np(bgk.get_div(vel, dx=dx, flags=flags))
    # collisions
    out["bgk"] = _np(bgk.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=force, mesh_grid=mesh))
    out["bgk_noforce"] = _np(bgk.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, mesh_grid=mesh))
    out["bgk_conc"] = _np(bgk.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=None,
                                        mesh_grid=mesh, is_convection=True, KBC_type=int(api.KBCType.KBC_C)))
    out["mrt"] = _np(mrt.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=force, mesh_grid=mesh,
                                   KBC_type=int(api.KBCType.KBC_C)))
    for name in ("LBGK", "KBC_A", "KBC_B", "KBC_C", "KBC_D"):
        out["kbc_" + name] = _np(kbc.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=force,
                                               mesh_grid=mesh, KBC_type=int(getattr(api.KBCType, name))))
    out["inputs_f"], out["inputs_flags"] = _np(f), _np(flags)
    return out
