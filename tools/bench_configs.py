#!/usr/bin/env python
"""Per-config throughput of the BASELINE.json configs other than the bench.py headline
(C1..C4 at native and bandwidth-bound sizes).  Not the driver's bench contract --
bench.py is -- but the same timing rules: warm-up, CUDA events (or synchronised wall clock around
whole scenario runs, differenced over two step counts so set-up cancels), inputs far larger than
L2 for the 8192^2 cases.

    python tools/bench_configs.py [--only c1,c2,...] [--big 8192]             (one GPU)
    torchrun --nproc-per-node N ... tools/bench_configs.py --only c4slab     (row slabs, NCCL)

Prints one JSON line per measurement:  {"config", "variant", "grid", "us_per_step", "mlups",
"gbs_72" (72 B per lattice update; 84 for the concentration step is reported as "gbs_alg")}.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies
sys.path.insert(1, ROOT)
sys.path.insert(2, os.path.join(ROOT, "tools"))

import torch  # noqa: E402


def emit(**kw):
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(kw), flush=True)


def scenario_us_per_step(S, api, dev, fn, kw, key, n1, n2):
    """(t(n2 steps) - t(n1 steps)) / (n2 - n1): per-step time of a whole scenario loop, host work
    of the demo included (the scenario functions mirror the demo loops line by line)."""
    def run(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        getattr(S, fn)(api, dev, **dict(kw, **{key: n}))
        torch.cuda.synchronize()
        return time.perf_counter() - t0
    run(n1)  # warm-up (allocator, module loads)
    t1, t2 = run(n1), run(n2)
    return (t2 - t1) / (n2 - n1) * 1e6


def fused_loop_us(step_fn, warm, K):
    for _ in range(warm):
        step_fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step_fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K * 1e3


def native_cases(S, api, dev):
    # C1 / C2 at the demos' 130 x 130 (demo_2d_LBM.py:144), C3 at 800 x 400 (demo_2d_LBM_cylinder.py:214),
    # C4 at 128 x 64 (demo_2d_LBM_JRL_C_cylinder.py), stage A at 128 x 64 (demo_2d_LBM_JRL_fluid_cylinder.py:201)
    cases = [
        ("c1", "shear_layer", dict(res=(130, 130), operator="bgk"), "steps", 130 * 130, (200, 1200), True),
        ("c2_mrt", "shear_layer", dict(res=(130, 130), operator="mrt", kbc_type=133), "steps", 130 * 130, (200, 1200), True),
        ("c2_kbc_a", "shear_layer", dict(res=(130, 130), operator="kbc", kbc_type=133), "steps", 130 * 130, (200, 1200), True),
        ("c3", "cylinder", dict(res=(800, 400), D=20), "steps", 800 * 400, (200, 1200), True),
        ("c4_conc_walk", "jrl_conc", dict(res=(128, 64), balance_steps=0), "walk_steps", 128 * 64, (100, 600), True),
        ("c5_fluid_stageA", "jrl_fluid", dict(res=(128, 64), gravity_strength=2e-5), "steps", 128 * 64, (100, 600), False),
    ]
    label = {False: "solver-object chain + demo host work", True: "fused step (1 launch per step from Python) + demo host work",
             "run": "jrl_run_f32: the whole time loop in one library call"}
    for name, fn, kw, key, cells, (n1, n2), has_fused in cases:
        modes = (False,) if not has_fused else ((False, True) if fn == "jrl_conc" else (False, True, "run"))
        for fused in modes:
            k2 = dict(kw, fused=fused) if fused else kw
            m1, m2 = (n1 * 10, n2 * 10) if fused == "run" else (n1, n2)
            us = scenario_us_per_step(S, api, dev, fn, k2, key, m1, m2)
            emit(config=name, variant=label[fused], grid=list(kw["res"]), us_per_step=round(us, 2), mlups=round(cells / us, 1))


def walk_cases(S, api, dev):
    """C4 native (128 x 64) walk loop with velocity snapshots interpolated as the demo does: host torch ops per step,
    device-side tip motion (1 launch), and the whole loop in one library call (jrl_conc_walk_f32)."""
    res = (128, 64)
    mesh = api.create_2d_meshgrid_tensor(size=(1, 1, *res), device=dev, dtype=torch.float32) + 0.5
    snaps = torch.cat([S.analytic_tip_velocity(mesh, 20.50000000000001 + k * 0.1, 6.50000000000001) for k in range(700)], dim=0).contiguous()
    kw = dict(res=res, balance_steps=0, vel_snaps=snaps, x_interval=0.1, fused=True)
    us = scenario_us_per_step(S, api, dev, "jrl_fluid", dict(res=res, gravity_strength=2e-5, device_tip="run"), "steps", 1000, 6000)
    emit(config="c5_fluid_stageA", variant="jrl_fluid_walk_f32: tip motion + stream + wall + moments + collide, whole loop in one library call",
         grid=list(res), us_per_step=round(us, 2), mlups=round(res[0] * res[1] / us, 1))
    for label, extra, (n1, n2) in (("fused step + host torch ops for the tip (demo loop)", {}, (100, 600)),
                                   ("fused step + device-side tip motion (jrl_tip_prepare_conc_f32), Python loop", dict(device_tip=True), (100, 600)),
                                   ("jrl_conc_walk_f32: tip motion + step + flux records, whole loop in one library call", dict(device_tip="run"), (1000, 6000))):
        us = scenario_us_per_step(S, api, dev, "jrl_conc", dict(kw, **extra), "walk_steps", n1, n2)
        emit(config="c4_conc_walk", variant=label, grid=list(res), us_per_step=round(us, 2), mlups=round(res[0] * res[1] / us, 1))


def big_cases(S, api, dev, G):
    import jrl_step
    import bench
    rho, vel = bench.shear_rows(G, torch.arange(G), dev)
    flags = torch.ones((1, 1, G, G), dtype=torch.uint8, device=dev)
    for op, kt in (("bgk", None), ("mrt", None), ("kbc", 133), ("kbc", 138)):
        P = api.SimulationParameters(simulation_size=(1, 1, G, G), device=dev, contact_angle=None, tau=bench.shear_tau(G))
        R = api.SimulationRunner(P)
        fs = R.create_fused_step(operator=op, kbc_type=kt)
        f = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
        st = [f, torch.empty_like(f)]

        def step():
            fs.step(st[0], flags, None, None, f_out=st[1], store_moments=False)
            st[0], st[1] = st[1], st[0]
        us = fused_loop_us(step, 3, 20)
        emit(config="c1" if op == "bgk" else "c2_" + op + ("" if kt is None else "_%d" % kt), variant="fused step, moments not stored",
             grid=[G, G], us_per_step=round(us, 1), mlups=round(G * G / us, 1), gbs_72=round(72.0 * G * G / us / 1e3, 1))
        del st, f
    torch.cuda.empty_cache()


def c4_state(api, S, dev, H, W, rows, refine):
    import workloads
    return workloads.c4_state(dev, H, W, rows, refine)


def c4_big(S, api, dev, G):
    import jrl_slab
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    refine = G // 128
    P = api.SimulationParameters(simulation_size=(1, 1, G, G), device=dev, contact_angle=None, axisymmetric_type=1)
    R = api.SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    row0, nrows = jrl_slab.slab_rows(G, world, rank)
    rows = torch.arange(row0 - 1, row0 + nrows + 1) % G if world > 1 else torch.arange(G)
    flags, mesh, phi, vel, C = c4_state(api, S, dev, G, G, rows, refine)
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    kw = dict(phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
    if world > 1:
        lat = jrl_slab.SlabLattice(fs, G, G, rank=rank, world=world, device=dev)
        lat.set_state(h, flags=flags, rho=C, vel=vel, **kw)
        Cout = torch.empty_like(C)

        def step():
            lat.step(store_moments=True, rho_out=Cout)
    else:
        st = [h, torch.empty_like(h), C, torch.empty_like(C)]

        def step():
            fs.step(st[0], flags, st[2], vel, f_out=st[1], rho_out=st[3], **kw)
            st[0], st[1], st[2], st[3] = st[1], st[0], st[3], st[2]
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 20
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    us = float(ms.item()) / K * 1e3
    # algorithmic bytes: h 9 reads + 9 writes, u 2 reads, C 1 write (84 B)
    emit(config="c4_conc_scaled", variant="fused concentration step, static tip, C stored every step", n_gpus=world,
         grid=[G, G], us_per_step=round(us, 1), mlups=round(G * G / us, 1), gbs_alg=round(84.0 * G * G / world / us / 1e3, 1),
         decomposition="row slabs (NCCL ring halo)" if world > 1 else "single domain")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="native,big,c4big")
    ap.add_argument("--big", type=int, default=8192)
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    import jrl_native
    jrl_native.load()
    import jrl_scenarios as S
    api = S.api_from_src()
    only = set(args.only.split(","))
    if "native" in only and world == 1:
        native_cases(S, api, dev)
    if ("native" in only or "walk" in only) and world == 1:
        walk_cases(S, api, dev)
    if "big" in only and world == 1:
        big_cases(S, api, dev, args.big)
    if "c4big" in only:
        c4_big(S, api, dev, args.big)
    if world > 1:
        import torch.distributed as dist
        import jrl_slab
        jrl_slab.release_all_graphs()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
