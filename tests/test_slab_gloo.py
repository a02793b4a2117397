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
