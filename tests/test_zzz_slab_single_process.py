"""The row-slab driver in ONE process (world = 1: the ring closes on itself, the ghost rows are the slab's own wrap rows)
against the plain single-domain step, bit for bit -- on the host emulation and, `-m gpu`, on the real library, where this
is the only pytest coverage of the slab FORM of the kernels (global-row mapping in jrl_step_f32 and jrl_collide_f32, halo
pack / unpack, edge rows and exchange on the comm stream beside the interior); the multi-rank runs are
tests/test_slab_gloo.py (CPU) and tools/check_slab_nccl.py (NCCL).

Collected after everything else on purpose (file name): JRL stage A on slabs (jrl_slab.AxisymFluidSlab) was built after
the round's GPU minutes were spent, so its GPU leg runs for the first time at the round's end and must not hide other
tests under `pytest -x`."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S
from conftest import have_cuda


def _bgk_problem(api, dev, kind):
    if kind == "periodic_wide":   # W >= 512: the pipelined kernel on the GPU
        H, W = 528, 1024   # > 2^19 cells: the interior rows take the pipelined kernel on the GPU
        flags = torch.ones((1, 1, H, W), dtype=torch.uint8, device=dev)
        rho, vel = S.shear_layer_state((H, W), dev)
        phi, inflow = None, None
    elif kind == "mixed_wide":
        H, W = 516, 1040
        st = S.random_state((H, W), 6, dev)
        flags, rho, vel, phi, inflow = st["flags"], st["rho"], st["vel"], st["phi"], 1.0
    else:
        H, W = 30, 24
        st = S.random_state((H, W), 5, dev)
        flags, rho, vel, phi, inflow = st["flags"], st["rho"], st["vel"], st["phi"], 1.0
    P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.7, contact_angle=None)
    R = api.SimulationRunner(P)
    f0 = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
    return R.create_fused_step(operator="bgk"), f0, flags, rho, vel, phi, inflow, (H, W)


def _check_bgk(api, dev, kind, steps=7):
    import jrl_slab
    fs, f0, flags, rho, vel, phi, inflow, (H, W) = _bgk_problem(api, dev, kind)
    f, r1, v1 = f0, rho, vel
    for _ in range(steps):
        f, r1, v1 = fs.step(f, flags, r1, v1, phi_obs=phi, inflow_density=inflow)
    lat = jrl_slab.SlabLattice(fs, H, W, rank=0, world=1, device=dev)
    fields = dict(flags=lat.scatter_global(flags))
    if phi is not None:
        fields.update(phi_obs=lat.scatter_global(phi), inflow_density=inflow)
    lat.set_state(lat.scatter_global(f0), **fields)
    r, v = lat.scatter_global(rho), lat.scatter_global(vel)
    for _ in range(steps):
        lat.step(rho=r, vel=v, store_moments=True)
        r, v = lat._staging["rho_out"].clone(), lat._staging["vel_out"].clone()
        lat.exchange(r)
        lat.exchange(v)
    for name, got, want in (("f", lat.f, f), ("rho", r, r1), ("vel", v, v1)):
        assert np.array_equal(lat.owned(got).cpu().numpy(), want.cpu().numpy(), equal_nan=True), f"{kind}: {name}"
    lat.close()


def _check_stage_a(api, dev, steps=7):
    import jrl_slab
    H, W = 36, 20
    st = S.random_state((H, W), 9, dev, axisym=1)
    P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.8, contact_angle=None, axisymmetric_type=1,
                                 gravity_strength=2e-5)
    R = api.SimulationRunner(P)
    col = R.create_collision()
    col.preset_KBC(dx=1.0, dt=1.0, tau=0.8)
    fs = R.create_fused_step(operator="none")
    mesh = api.create_2d_meshgrid_tensor(size=(1, 1, H, W), device=dev, dtype=torch.float32) + 0.5
    op, kbc, tau = col._operator(False, None)
    f, r1, v1 = st["f"], st["rho"], st["vel"]
    for _ in range(steps):
        f1, r1, v1 = fs.step(f, st["flags"], r1, v1, phi_obs=st["phi"], use_wall=True, obs_vel=st["obs_vel"], inflow_density=1.0)
        f = col._collide(1.0, 1.0, f1, r1, v1, st["flags"], st["force"], mesh, False, op, kbc, tau)
    lat = jrl_slab.SlabLattice(fs, H, W, rank=0, world=1, device=dev)
    lat.set_state(lat.scatter_global(st["f"]), flags=lat.scatter_global(st["flags"]), phi_obs=lat.scatter_global(st["phi"]),
                  use_wall=True, obs_vel=lat.scatter_global(st["obs_vel"]), inflow_density=1.0)
    stage = jrl_slab.AxisymFluidSlab(lat, col)
    r, v = lat.scatter_global(st["rho"]), lat.scatter_global(st["vel"])
    mesh_s, force_s = lat.scatter_global(mesh), lat.scatter_global(st["force"])
    for _ in range(steps):
        r, v = stage.step(r, v, mesh_s, force=force_s)
    for name, got, want in (("f", lat.f, f), ("rho", r, r1), ("vel", v, v1)):
        assert np.array_equal(lat.owned(got).cpu().numpy(), want.cpu().numpy(), equal_nan=True), f"stage A: {name}"
    lat.close()


KINDS = ["periodic_wide", "mixed_wide", "mixed"]


@pytest.mark.skipif(have_cuda(), reason="host emulation is for the CPU-only container")
@pytest.mark.parametrize("kind", KINDS)
def test_slab_form_equals_single_domain_emu(emu_api, kind):
    _check_bgk(emu_api, torch.device("cpu"), kind)


@pytest.mark.skipif(have_cuda(), reason="host emulation is for the CPU-only container")
def test_stage_a_slab_form_equals_single_domain_emu(emu_api):
    _check_stage_a(emu_api, torch.device("cpu"))


@pytest.mark.gpu
@pytest.mark.parametrize("kind", KINDS)
def test_slab_form_equals_single_domain_gpu(cuda_api, kind):
    _check_bgk(cuda_api, torch.device("cuda:0"), kind)


@pytest.mark.gpu
def test_stage_a_slab_form_equals_single_domain_gpu(cuda_api):
    _check_stage_a(cuda_api, torch.device("cuda:0"))
