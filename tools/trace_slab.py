#!/usr/bin/env python
"""Device timeline of the row-slab concentration step (the C4 scaled configuration of bench.py) over NCCL, taken with
torch.profiler's CUDA activity records (kineto; there is no nsys on the box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
        tools/trace_slab.py [--rows-per-rank 1024] [--width 8192] [--steps 6]

Every rank advances `rows-per-rank` rows of a (rows-per-rank x world) x width lattice -- with 1024 x 8192 the per-rank
work of the 8192^2 grid on 8 GPUs, reproduced on 2.  Rank 0 writes gpurun_out/slab_trace.json: for the profiled steps
every kernel / memcpy / memset with its stream, start (us, relative to the first record) and duration, plus the
CUDA-event step time WITHOUT the profiler (the profiler's own numbers are for the order of things, not for speed)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-per-rank", type=int, default=1024)
    ap.add_argument("--width", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--workload", choices=("c4", "bgk"), default="c4", help="c4: concentration step of bench.py's c4_conc_scaled; "
                    "bgk: the headline's periodic all-FLUID BGK step")
    ap.add_argument("--no-trace", action="store_true", help="CUDA-event timings only")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "slab_trace.json"))
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import jrl_native
    jrl_native.load()
    import jrl_slab
    import workloads as WL
    from src.LBM.simulation import SimulationParameters, SimulationRunner
    H, W = args.rows_per_rank * world, args.width
    row0, nrows = jrl_slab.slab_rows(H, world, rank)
    rows = torch.arange(row0 - 1, row0 + nrows + 1) % H
    if args.workload == "c4":
        P = SimulationParameters(simulation_size=(1, 1, H, W), device=dev, contact_angle=None, axisymmetric_type=1)
        R = SimulationRunner(P)
        fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
        flags, mesh, phi, vel, C = WL.c4_state(dev, H, W, rows, max(1, min(H, W) // 128))
        h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
        lat = jrl_slab.SlabLattice(fs, H, W, rank=rank, world=world, device=dev)
        lat.set_state(h, flags=flags, rho=C, vel=vel, phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
        Cout = torch.empty_like(C)

        def step():
            lat.step(store_moments=True, rho_out=Cout)
    else:
        P = SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.7, contact_angle=None)
        R = SimulationRunner(P)
        fs = R.create_fused_step(operator="bgk")
        flags = torch.ones((1, 1, nrows + 2, W), dtype=torch.uint8, device=dev)
        y = (rows.to(dev).float().view(1, 1, -1, 1) + 0.5) / H
        x = (torch.arange(W, device=dev).float().view(1, 1, 1, -1) + 0.5) / W
        rho = torch.ones((1, 1, nrows + 2, W), device=dev)
        vel = torch.cat([0.04 * torch.tanh(80.0 * (0.25 - (y - 0.5).abs())).expand(1, 1, nrows + 2, W),
                         (0.002 * torch.sin(6.2831853 * (x + 0.25))).expand(1, 1, nrows + 2, W)], dim=1).contiguous()
        f0 = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)
        del rho, vel
        lat = jrl_slab.SlabLattice(fs, H, W, rank=rank, world=world, device=dev)
        lat.set_state(f0, flags=flags)

        def step():
            lat.step()

    def timed(n):
        for _ in range(5):
            step()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    report = {"world": world, "workload": args.workload, "rows_per_rank": nrows, "width": W, "order": os.environ.get("JRL_SLAB_ORDER", "interior_first"),
              "graph": os.environ.get("JRL_SLAB_GRAPH", "0"), "ms_per_step_overlap": timed(40)}
    if args.no_trace:
        if rank == 0:
            print(json.dumps(report), flush=True)
        dist.barrier()
        jrl_slab.release_all_graphs()
        dist.destroy_process_group()
        return
    lat.overlap = False
    report["ms_per_step_no_overlap"] = timed(40)
    lat.overlap = True
    # interior alone / exchange alone, for the arithmetic of what could hide behind what
    ex = lat.exchange

    lat.exchange = lambda field: None
    report["ms_per_step_no_exchange"] = timed(40)
    lat.exchange = ex

    def only_exchange():
        ex(lat.f)
    for _ in range(5):
        only_exchange()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(40):
        only_exchange()
    e1.record()
    torch.cuda.synchronize()
    report["ms_per_exchange_alone"] = e0.elapsed_time(e1) / 40

    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize()
    dist.barrier()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            step()
        torch.cuda.synchronize()
    # the stream of a record is not on FunctionEvent: take it from the chrome trace instead
    tmp = args.out + f".chrome.{rank}.json"
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    prof.export_chrome_trace(tmp)
    with open(tmp) as fh:
        tr = json.load(fh)
    os.remove(tmp)
    gpu = [e for e in tr.get("traceEvents", []) if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    gpu.sort(key=lambda e: e["ts"])
    t0 = gpu[0]["ts"] if gpu else 0
    report["timeline"] = [dict(t_us=round(e["ts"] - t0, 1), dur_us=round(e["dur"], 1), stream=e.get("args", {}).get("stream"),
                               name=e["name"][:70]) for e in gpu]
    report["span_us_per_step"] = (gpu[-1]["ts"] + gpu[-1]["dur"] - t0) / args.steps if gpu else None
    if rank == 0:
        with open(args.out, "w") as fh:
            json.dump(report, fh, indent=1)
        print(json.dumps({k: v for k, v in report.items() if k != "timeline"}), flush=True)
        for e in report["timeline"][: 14 * 3]:
            print(f'{e["t_us"]:9.1f} {e["dur_us"]:8.1f}  s{e["stream"]}  {e["name"]}')
    dist.barrier()
    jrl_slab.release_all_graphs()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
