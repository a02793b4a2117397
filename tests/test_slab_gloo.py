"""CPU, world_size 2 and 3 over gloo: the row-slab driver (jrl_slab.SlabLattice -- halo exchange,
edge/interior split, ring wrap) reproduces the single-domain result bit for bit.  Kernels run
through the host-emulation build; the GPU box runs the same driver over NCCL (bench.py --gpus N)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, build_emu, have_cuda

H, W, STEPS = 30, 24, 9


def _setup():
    sys.path.insert(0, PKG)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emu"))
    import install as emu_install
    emu_install.install(build_emu())
    import jrl_scenarios as S
    return S, S.api_from_src()


def _problem(api, S, kind):
    dev = torch.device("cpu")
    g = torch.Generator().manual_seed(11)
    if kind == "periodic":  # C1: all FLUID, periodic in both directions (ring wrap crosses ranks)
        flags = torch.ones((1, 1, H, W), dtype=torch.uint8)
        rho, vel = S.shear_layer_state((H, W), dev)
        phi = None
    else:                   # obstacles / inlets / outlets straddling the cuts, interpolated bounce-back
        st = S.random_state((H, W), 5, dev)
        flags, rho, vel, phi = st["flags"], st["rho"], st["vel"], st["phi"]
    P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.7, contact_angle=None)
    R = api.SimulationRunner(P)
    f0 = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
    fs = R.create_fused_step(operator="bgk")
    return fs, f0, flags, rho, vel, phi


def _conc_problem(api):
    """BASELINE config 4's concentration step (tools/workloads.c4_state: probe tip, x-axis symmetry, INFLOW ring, OUTFLOW
    edges, interpolated bounce-back, clamp) at 48 x 32; the tip's solid block and its inflow row straddle the slab cuts."""
    sys.path.insert(0, os.path.join(os.path.dirname(PKG), "tools"))
    import workloads as WL
    dev = torch.device("cpu")
    Hc, Wc = 48, 32
    P = api.SimulationParameters(simulation_size=(1, 1, Hc, Wc), device=dev, contact_angle=None, axisymmetric_type=1)
    R = api.SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    flags, mesh, phi, vel, C = WL.c4_state(dev, Hc, Wc, torch.arange(Hc), 1)
    C = (C * (0.6 + 0.4 * torch.rand(C.shape, generator=torch.Generator().manual_seed(2)))).contiguous()
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    return fs, h, flags, mesh, phi, vel, C, (Hc, Wc)


def _single_conc():
    S, api = _setup()
    fs, h, flags, mesh, phi, vel, C, _ = _conc_problem(api)
    for _ in range(STEPS):
        h, C = fs.step(h, flags, C, vel, phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
    return h.numpy(), C.numpy()


def _worker_conc(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S, api = _setup()
    import jrl_slab
    fs, h, flags, mesh, phi, vel, C, (Hc, Wc) = _conc_problem(api)
    lat = jrl_slab.SlabLattice(fs, Hc, Wc, device="cpu")
    lat.set_state(lat.scatter_global(h), flags=lat.scatter_global(flags), vel=lat.scatter_global(vel),
                  phi_obs=lat.scatter_global(phi), mesh_grid=lat.scatter_global(mesh), inflow_density=0, clamp01=True)
    c = lat.scatter_global(C)
    for _ in range(STEPS):
        lat.step(rho=c, store_moments=True)
        c = lat._staging["rho_out"].clone()
        lat.exchange(c)
    np.save(os.path.join(out_dir, f"h_{rank}.npy"), lat.owned(lat.f).numpy())
    np.save(os.path.join(out_dir, f"C_{rank}.npy"), lat.owned(c).numpy())
    dist.barrier()
    dist.destroy_process_group()


def _wall_problem(api):
    """BASELINE config 3 (tools/workloads.cylinder_state: MRT, INFLOW / OUTFLOW rows, moving-wall reconstruction around a
    cylinder) at 48 x 36 with the cylinder centred on a slab cut."""
    sys.path.insert(0, os.path.join(os.path.dirname(PKG), "tools"))
    import workloads as WL
    dev = torch.device("cpu")
    Hc, Wc = 48, 36
    flags, phi, rho, vel, inflow_vel = WL.cylinder_state(dev, Hc, Wc, 9, (17.0, 24.5))
    P = api.SimulationParameters(simulation_size=(1, 1, Hc, Wc), device=dev, tau=0.56, contact_angle=None)
    R = api.SimulationRunner(P)
    fs = R.create_fused_step(operator="mrt")
    f0 = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
    return fs, f0, flags, phi, rho, vel, inflow_vel, (Hc, Wc)


def _single_wall():
    S, api = _setup()
    fs, f, flags, phi, rho, vel, inflow_vel, _ = _wall_problem(api)
    zero = torch.zeros_like(vel)
    for _ in range(STEPS):
        f, rho, vel = fs.step(f, flags, rho, vel, inflow_density=1.0, inflow_vel=inflow_vel, use_wall=True, wall_phi=phi, obs_vel=zero)
    return f.numpy(), rho.numpy(), vel.numpy()


def _worker_wall(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S, api = _setup()
    import jrl_slab
    fs, f0, flags, phi, rho, vel, inflow_vel, (Hc, Wc) = _wall_problem(api)
    lat = jrl_slab.SlabLattice(fs, Hc, Wc, device="cpu")
    lat.set_state(lat.scatter_global(f0), flags=lat.scatter_global(flags), inflow_density=1.0, inflow_vel=inflow_vel,
                  use_wall=True, wall_phi=lat.scatter_global(phi), obs_vel=lat.scatter_global(torch.zeros_like(vel)))
    r, v = lat.scatter_global(rho), lat.scatter_global(vel)
    for _ in range(STEPS):
        lat.step(rho=r, vel=v, store_moments=True)
        r, v = lat._staging["rho_out"].clone(), lat._staging["vel_out"].clone()
        lat.exchange(r)
        lat.exchange(v)
    for name, t in (("f", lat.f), ("rho", r), ("vel", v)):
        np.save(os.path.join(out_dir, f"{name}_{rank}.npy"), lat.owned(t).numpy())
    dist.barrier()
    dist.destroy_process_group()


def _stage_a_problem(api, S):
    """JRL stage A's operator chain (x-axis symmetry, interpolated bounce-back, moving wall, gravity, BGK with the
    axisymmetric source) on a seeded random state holding every CellType -- obstacles, inlets and outlets on the cuts."""
    dev = torch.device("cpu")
    Hc, Wc = 36, 20
    st = S.random_state((Hc, Wc), 9, dev, axisym=1)
    P = api.SimulationParameters(simulation_size=(1, 1, Hc, Wc), device=dev, tau=0.8, contact_angle=None, axisymmetric_type=1,
                                 gravity_strength=2e-5)
    R = api.SimulationRunner(P)
    col = R.create_collision()
    col.preset_KBC(dx=1.0, dt=1.0, tau=0.8)
    fs = R.create_fused_step(operator="none")
    mesh = api.create_2d_meshgrid_tensor(size=(1, 1, Hc, Wc), device=dev, dtype=torch.float32) + 0.5
    return fs, col, st, mesh, (Hc, Wc)


def _single_stage_a():
    S, api = _setup()
    fs, col, st, mesh, _ = _stage_a_problem(api, S)
    f, rho, vel = st["f"], st["rho"], st["vel"]
    op, kbc, tau = col._operator(False, None)
    for _ in range(STEPS):
        f1, rho, vel = fs.step(f, st["flags"], rho, vel, phi_obs=st["phi"], use_wall=True, obs_vel=st["obs_vel"], inflow_density=1.0)
        f = col._collide(1.0, 1.0, f1, rho, vel, st["flags"], st["force"], mesh, False, op, kbc, tau)
    return f.numpy(), rho.numpy(), vel.numpy()


def _worker_stage_a(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S, api = _setup()
    import jrl_slab
    fs, col, st, mesh, (Hc, Wc) = _stage_a_problem(api, S)
    lat = jrl_slab.SlabLattice(fs, Hc, Wc, device="cpu")
    lat.set_state(lat.scatter_global(st["f"]), flags=lat.scatter_global(st["flags"]), phi_obs=lat.scatter_global(st["phi"]),
                  use_wall=True, obs_vel=lat.scatter_global(st["obs_vel"]), inflow_density=1.0)
    stage = jrl_slab.AxisymFluidSlab(lat, col)
    r, v = lat.scatter_global(st["rho"]), lat.scatter_global(st["vel"])
    mesh_s, force_s = lat.scatter_global(mesh), lat.scatter_global(st["force"])
    for _ in range(STEPS):
        r, v = stage.step(r, v, mesh_s, force=force_s)
    for name, t in (("f", lat.f), ("rho", r), ("vel", v)):
        np.save(os.path.join(out_dir, f"{name}_{rank}.npy"), lat.owned(t).numpy())
    dist.barrier()
    dist.destroy_process_group()


def _single(kind):
    S, api = _setup()
    fs, f, flags, rho, vel, phi = _problem(api, S, kind)
    for _ in range(STEPS):
        f, rho, vel = fs.step(f, flags, rho, vel, phi_obs=phi, inflow_density=1.0 if kind != "periodic" else None)
    return f.numpy()


def _worker(rank, world, port, kind, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S, api = _setup()
    import jrl_slab
    fs, f0, flags, rho, vel, phi = _problem(api, S, kind)
    lat = jrl_slab.SlabLattice(fs, H, W, device="cpu")
    fields = dict(flags=lat.scatter_global(flags))
    if kind != "periodic":
        # previous-step moments are inputs at OBSTACLE / INFLOW cells: keep them current per step
        fields.update(phi_obs=lat.scatter_global(phi), inflow_density=1.0)
    lat.set_state(lat.scatter_global(f0), **fields)
    r, v = lat.scatter_global(rho), lat.scatter_global(vel)
    for _ in range(STEPS):
        if kind == "periodic":
            lat.step()
        else:
            lat.step(rho=r, vel=v, store_moments=True)
            r, v = lat._staging["rho_out"].clone(), lat._staging["vel_out"].clone()
            lat.exchange(r)
            lat.exchange(v)
    np.save(os.path.join(out_dir, f"slab_{rank}.npy"), lat.owned(lat.f).numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(have_cuda(), reason="CPU/gloo test of the host-side logic")
@pytest.mark.parametrize("world,kind", [(2, "periodic"), (3, "periodic"), (2, "mixed")])
def test_slab_equals_single_domain(tmp_path, world, kind):
    build_emu()
    port = 29500 + (os.getpid() + world * 7 + len(kind)) % 2000
    mp.spawn(_worker, args=(world, port, kind, str(tmp_path)), nprocs=world, join=True)
    got = np.concatenate([np.load(tmp_path / f"slab_{r}.npy") for r in range(world)], axis=2)
    want = _single(kind)
    assert np.array_equal(got, want, equal_nan=True)


@pytest.mark.skipif(have_cuda(), reason="CPU/gloo test of the host-side logic")
@pytest.mark.parametrize("world", [3])
def test_concentration_slab_equals_single_domain(tmp_path, world):
    """BASELINE config 4 is "row-slab sharded": the fused concentration step on slabs == the single domain, bit for bit."""
    build_emu()
    port = 31500 + (os.getpid() + world * 11) % 2000
    mp.spawn(_worker_conc, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    h_want, C_want = _single_conc()
    h = np.concatenate([np.load(tmp_path / f"h_{r}.npy") for r in range(world)], axis=2)
    C = np.concatenate([np.load(tmp_path / f"C_{r}.npy") for r in range(world)], axis=2)
    assert np.array_equal(h, h_want, equal_nan=True)
    assert np.array_equal(C, C_want, equal_nan=True)
    assert float(np.abs(C_want - 1.0).max()) > 1e-3   # the field is not trivially uniform


@pytest.mark.skipif(have_cuda(), reason="CPU/gloo test of the host-side logic")
@pytest.mark.parametrize("world", [2])
def test_moving_wall_slab_equals_single_domain(tmp_path, world):
    """MRT + moving-wall reconstruction + INFLOW / OUTFLOW rows (the C3 step) on slabs == the single domain, bit for bit:
    the wall step reads the pre-step velocity of neighbour rows, so rho / vel ghost rows are exchanged per step too."""
    build_emu()
    port = 33500 + (os.getpid() + world * 13) % 2000
    mp.spawn(_worker_wall, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    want = dict(zip(("f", "rho", "vel"), _single_wall()))
    for name, w in want.items():
        got = np.concatenate([np.load(tmp_path / f"{name}_{r}.npy") for r in range(world)], axis=2)
        assert np.array_equal(got, w, equal_nan=True), name
    assert float(np.abs(want["vel"]).max()) > 1e-3


@pytest.mark.skipif(have_cuda(), reason="CPU/gloo test of the host-side logic")
@pytest.mark.parametrize("world", [2, 3])
def test_stage_a_slab_equals_single_domain(tmp_path, world):
    """JRL stage A (axisymmetric fluid: fused first pass, moments exchanged, collision with the stencil source in its slab
    form) on slabs == the single domain's two launches, bit for bit (jrl_slab.AxisymFluidSlab)."""
    build_emu()
    port = 35500 + (os.getpid() + world * 17) % 2000
    mp.spawn(_worker_stage_a, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    want = dict(zip(("f", "rho", "vel"), _single_stage_a()))
    for name, w in want.items():
        got = np.concatenate([np.load(tmp_path / f"{name}_{r}.npy") for r in range(world)], axis=2)
        assert np.array_equal(got, w, equal_nan=True), name
    assert np.isfinite(want["f"]).mean() > 0.9
