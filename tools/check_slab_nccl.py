#!/usr/bin/env python
"""Multi-GPU parity check of the row-slab driver over NCCL (run under torchrun on a GPU box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/check_slab_nccl.py [--grid 2048] [--steps 12]

Every rank advances (a) its slab of the distributed lattice (jrl_slab.SlabLattice: edge rows ->
halo exchange overlapped with the interior) and (b) the whole lattice on its own GPU as a single
domain; the owned rows of (a) must equal the same rows of (b) BIT FOR BIT.  Two problems: the
periodic all-FLUID shear layer (ring wrap crosses ranks; pipelined TMA kernel) and a mixed grid
with obstacles / inlets / outlets / interpolated bounce-back straddling the cuts.  Prints one
JSON line on rank 0; exit code 1 on any mismatch.  The CPU suite runs the same driver over gloo
(tests/test_slab_gloo.py)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--stage-a", action="store_true", help="also check jrl_slab.AxisymFluidSlab (not yet run on GPUs)")
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    if os.environ.get("JRL_DEBUG_HANG"):  # print every thread's stack and exit if the check has not finished by then
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["JRL_DEBUG_HANG"]), exit=True)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import jrl_native
    jrl_native.load()
    import jrl_scenarios as S
    import jrl_slab
    api = S.api_from_src()
    H = W = args.grid
    report = {"world": world, "grid": args.grid, "steps": args.steps}
    bad = 0
    for kind in ("periodic", "mixed"):
        if kind == "periodic":
            flags = torch.ones((1, 1, H, W), dtype=torch.uint8, device=dev)
            rho, vel = S.shear_layer_state((H, W), dev)
            phi = None
        else:
            st = S.random_state((H, W), 5, dev)
            flags, rho, vel, phi = st["flags"], st["rho"], st["vel"], st["phi"]
        P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.7, contact_angle=None)
        R = api.SimulationRunner(P)
        f0 = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
        fs = R.create_fused_step(operator="bgk")
        # (b) single domain on this GPU
        f, r1, v1 = f0, rho, vel
        for _ in range(args.steps):
            f, r1, v1 = fs.step(f, flags, r1, v1, phi_obs=phi, inflow_density=1.0 if kind != "periodic" else None)
        # (a) slabs
        lat = jrl_slab.SlabLattice(fs, H, W, device=dev)
        fields = dict(flags=lat.scatter_global(flags))
        if kind != "periodic":
            fields.update(phi_obs=lat.scatter_global(phi), inflow_density=1.0)
        lat.set_state(lat.scatter_global(f0), **fields)
        r, v = lat.scatter_global(rho), lat.scatter_global(vel)
        for _ in range(args.steps):
            if kind == "periodic":
                lat.step()
            else:
                lat.step(rho=r, vel=v, store_moments=True)
                r, v = lat._staging["rho_out"].clone(), lat._staging["vel_out"].clone()
                lat.exchange(r)
                lat.exchange(v)
        torch.cuda.synchronize()
        mine = lat.owned(lat.f)
        want = f[:, :, lat.row0:lat.row0 + lat.nrows, :]
        same = torch.equal(mine, want) or bool(((mine == want) | (mine.isnan() & want.isnan())).all())
        t = torch.tensor([0 if same else 1], device=dev)
        dist.all_reduce(t)
        report[kind] = "bit-exact" if t.item() == 0 else f"MISMATCH on {int(t.item())} rank(s)"
        bad += int(t.item())
    # JRL stage A's two-pass step on slabs (jrl_slab.AxisymFluidSlab) against the single domain's two launches: the
    # problem of tests/test_slab_gloo.py (seeded random state with every CellType, x-axis symmetry, moving wall, gravity)
    if args.stage_a:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        Hc, Wc = 36 * world, 2048
        st = S.random_state((Hc, Wc), 9, dev, axisym=1)
        P = api.SimulationParameters(simulation_size=(1, 1, Hc, Wc), device=dev, tau=0.8, contact_angle=None, axisymmetric_type=1,
                                     gravity_strength=2e-5)
        R = api.SimulationRunner(P)
        col = R.create_collision()
        col.preset_KBC(dx=1.0, dt=1.0, tau=0.8)
        fs = R.create_fused_step(operator="none")
        mesh = api.create_2d_meshgrid_tensor(size=(1, 1, Hc, Wc), device=dev, dtype=torch.float32) + 0.5
        op, kbc, tau = col._operator(False, None)
        f, r1, v1 = st["f"], st["rho"], st["vel"]
        for _ in range(args.steps):
            f1, r1, v1 = fs.step(f, st["flags"], r1, v1, phi_obs=st["phi"], use_wall=True, obs_vel=st["obs_vel"], inflow_density=1.0)
            f = col._collide(1.0, 1.0, f1, r1, v1, st["flags"], st["force"], mesh, False, op, kbc, tau)
        lat = jrl_slab.SlabLattice(fs, Hc, Wc, device=dev)
        lat.set_state(lat.scatter_global(st["f"]), flags=lat.scatter_global(st["flags"]), phi_obs=lat.scatter_global(st["phi"]),
                      use_wall=True, obs_vel=lat.scatter_global(st["obs_vel"]), inflow_density=1.0)
        stage = jrl_slab.AxisymFluidSlab(lat, col)
        r, v = lat.scatter_global(st["rho"]), lat.scatter_global(st["vel"])
        mesh_s, force_s = lat.scatter_global(mesh), lat.scatter_global(st["force"])
        for _ in range(args.steps):
            r, v = stage.step(r, v, mesh_s, force=force_s)
        torch.cuda.synchronize()
        mine, want = lat.owned(lat.f), f[:, :, lat.row0:lat.row0 + lat.nrows, :]
        same = bool(((mine == want) | (mine.isnan() & want.isnan())).all())
        t = torch.tensor([0 if same else 1], device=dev)
        dist.all_reduce(t)
        report["stage_a"] = "bit-exact" if t.item() == 0 else f"MISMATCH on {int(t.item())} rank(s)"
        bad += int(t.item())
    if rank == 0:
        print(json.dumps(report), flush=True)
    dist.barrier()
    jrl_slab.release_all_graphs()
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
