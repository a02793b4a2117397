"""Test / baseline infrastructure (oracle/): the reference's per-step loop on the CPU, as fast as
the reference itself could run it there.  NOT product code; only bench.py's `cpu_baseline` leg and
`--impl reference` arm and tests/ may import it.

What runs:
  * streaming K1  = the reference's OWN kernel body (propagation.cu:22-488) compiled for the host
    with OpenMP by oracle/build_ref.sh (oracle/_ref/libjrl_ref.so), called as the reference's
    wrapper does, including its `f_new = f + 0.0` copy (LBM_propagation_2d.py:49);
  * moments, equilibrium and BGK collision = the reference's torch-op chains restated with the
    same torch operators on CPU tensors (they ARE torch ops in the reference:
    LBM_macro_compute_2d.py:316-351, LBM_collision_2d.py:236-263, 473-555), so torch's intra-op
    thread pool uses every host core exactly as it would under the reference.
The reference's Python package itself cannot travel to the GPU box (/root/reference is absent
there) and hard-requires CUDA, hence this arrangement.  tests/test_oracle_golden.py
checks this module against the golden fixtures written by the real reference classes.
"""
from __future__ import annotations

import numpy as np
import torch

import ref_cpu

_E = torch.tensor([[0, 0], [1, 0], [0, 1], [-1, 0], [0, -1], [1, 1], [-1, 1], [-1, -1], [1, -1]],
                  dtype=torch.int64).reshape(1, 9, 2, 1, 1)
_W = torch.tensor([4 / 9] + [1 / 9] * 4 + [1 / 36] * 4, dtype=torch.float32).reshape(1, 9, 1, 1)
OBSTACLE = 2


def get_feq(dx, dt, rho, vel, force=None):
    """LBM_collision_2d.py:236-263."""
    if force is not None:
        vel = vel + torch.where(rho <= 1e-10, torch.zeros_like(vel), 0.5 * force / rho)
    c = dx / dt
    cs2 = c * c / 3.0
    eu = (vel.unsqueeze(1) * _E * c).sum(dim=2)
    feq = rho * _W * (1.0 + eu / cs2)
    uv = (vel * vel).sum(dim=1, keepdim=True)
    feq += rho * _W * (0.5 * eu * eu / cs2 / cs2 - 0.5 * uv / cs2)
    return feq


def macro_compute(dx, dt, f, rho, vel, flags):
    """LBM_macro_compute_2d.py:316-351 with contact_angle=None."""
    c = dx / dt
    obs = flags == OBSTACLE
    rho_new = torch.where(obs, rho, f.sum(dim=1).unsqueeze(1))
    vel_new = (f.unsqueeze(2) * _E).sum(dim=1) * (c / rho_new)
    return rho_new, torch.where(obs, vel, vel_new)


def collide_bgk(dx, dt, f, rho, vel, flags, tau, force=None):
    """LBM_collision_2d.py:473-555, non-axisymmetric fluid branch (Gi = 0.0)."""
    feq = get_feq(dx, dt, rho, vel, force)
    out = f + (1.0 / tau) * (feq - f) + dt * 0.0
    return torch.where(flags == OBSTACLE, f, out)


class ReferenceCpuLoop:
    """demo_2d_LBM.py:101-119 loop body on CPU tensors."""

    def __init__(self, tau, threads=None):
        self.tau = tau
        self.ker = ref_cpu.RefKernels()
        if threads:
            torch.set_num_threads(threads)
            self.ker.set_num_threads(threads)   # the OpenMP runtime of libjrl_ref.so (OMP_NUM_THREADS may be 1 under torchrun)

    @property
    def cores(self):
        return max(torch.get_num_threads(), self.ker.num_threads)

    def step(self, f, rho, vel, flags, force=None):
        f_new = torch.from_numpy(self.ker.prop(flags.numpy(), f.numpy()))     # K1 (+ the wrapper's copy)
        rho, vel = macro_compute(1.0, 1.0, f_new, rho, vel, flags)
        f = collide_bgk(1.0, 1.0, f_new, rho, vel, flags, self.tau, force)
        return f, rho, vel
