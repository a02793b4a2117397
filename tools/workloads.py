"""Synthetic workload builders for the bandwidth-bound (scaled) BASELINE configs, shared by bench.py's `configs` leg and
tools/bench_configs.py.  Geometry and parameters follow the demos they scale (cited per function); everything is built
on the device, row range by row range, so a row slab never materialises the whole grid.  No oracle / test code here."""
import math

import torch


def analytic_tip_velocity(mesh_grid, walks, r0):
    """Stand-in for the stage-A velocity field (SURVEY 8d, C4 option (b)): u = (-x / (2 r0), 1) exp(-d / r0), d = distance
    of the cell centre from the tip-face centre, normalised to the tip speed."""
    x, y = mesh_grid[:, 0:1], mesh_grid[:, 1:2]
    w = torch.exp(-torch.sqrt(x * x + (y - walks) * (y - walks)) / r0)
    return torch.cat([-x / (2.0 * r0) * w, w], dim=1)


def c4_state(dev, H, W, rows, refine):
    """Scaled C4: tip radius 6.5 x refine, inflow height 20.5 x refine, Rg = 4, x-axis symmetry, is_negative
    layout (demo_2d_LBM_JRL_C_cylinder.py:151-161 with the `refine` semantics of demo_2d_LBM_JRL_C_ethaline.py:79-89), tip at
    its start position, analytic stand-in velocity.  `rows`: the global rows to build (wrapped LongTensor).
    Returns flags [1,1,R,W] u8, mesh [1,2,R,W], phi [1,1,R,W], vel [1,2,R,W], C [1,1,R,W]."""
    radius_obs = 6.0 * refine + 0.5
    inflow_height = 20.5 * refine
    Rg_radius = int(4 * radius_obs + 0.5)
    hi, ri = int(inflow_height + 0.5), int(radius_obs + 0.5)
    y = rows.to(dev).view(-1, 1)
    x = torch.arange(W, device=dev).view(1, -1)
    fl = torch.full((rows.numel(), W), 1, dtype=torch.uint8, device=dev)
    fl[(y == 0).expand_as(fl) & (x >= 1)] = 16
    fl[(y == H - 1).expand_as(fl)] = 2
    fl[(x == W - 1).expand_as(fl) & (y < H - 1)] = 32
    fl[(y == 0).expand_as(fl) & (x >= Rg_radius)] = 32
    fl[(y < hi) & (x < Rg_radius)] = 2
    fl[(y == hi - 1) & (x < ri)] = 8
    mesh = torch.stack([(x + 0.5).expand(rows.numel(), W), (y + 0.5).expand(rows.numel(), W)]).unsqueeze(0).float().contiguous()
    box_radius = Rg_radius + 100.0
    phi = torch.maximum(mesh[:, 0:1] - (Rg_radius - box_radius) - box_radius, mesh[:, 1:2] - (inflow_height - box_radius) - box_radius)
    phi = torch.minimum(phi, -(mesh[:, 1:2] - H)).contiguous()
    vel = (analytic_tip_velocity(mesh, inflow_height, radius_obs) * 0.05).contiguous()
    C = torch.ones((1, 1, rows.numel(), W), device=dev)
    C[0, 0][fl == 2] = 0.0
    return fl.view(1, 1, -1, W).contiguous(), mesh, phi, vel, C


def cylinder_state(dev, H, W, D, center, Vmax=0.05):
    """demo_2d_LBM_cylinder.py:111-136 at any size: OBSTACLE ring, INFLOW row 0, OUTFLOW row H - 1, cylinder of diameter D
    from the signed distance at cell centres.  Returns flags, phi_obs, rho, vel, inflow_vel."""
    y = torch.arange(H, device=dev, dtype=torch.float32).view(H, 1) + 0.5
    x = torch.arange(W, device=dev, dtype=torch.float32).view(1, W) + 0.5
    flags = torch.full((1, 1, H, W), 2, dtype=torch.uint8, device=dev)
    flags[..., 1:-1, 1:-1] = 1
    flags[..., 0, 1:-1] = 8
    flags[..., -1, 1:-1] = 16
    d2 = (x - center[0]) * (x - center[0]) + (y - center[1]) * (y - center[1]) - (D / 2.0) * (D / 2.0)
    phi = torch.where(d2 > 0, torch.sqrt(d2.clamp_min(0)), -torch.sqrt((-d2).clamp_min(0))).view(1, 1, H, W).contiguous()
    flags[phi <= 0] = 2
    inflow_vel = torch.tensor([0.0, Vmax], dtype=torch.float32, device=dev)
    rho = torch.ones((1, 1, H, W), dtype=torch.float32, device=dev)
    vel = torch.zeros((1, 2, H, W), dtype=torch.float32, device=dev)
    vel[:, 1][flags[:, 0] == 8] = Vmax
    return flags, phi, rho, vel, inflow_vel


def cylinder_tau(H, W, Vmax=0.05, Re=5000.0):
    return 0.5 + (Vmax * max(H, W) / Re) / (1.0 / 3.0)   # demo_2d_LBM_cylinder.py:52-60
