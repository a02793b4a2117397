#!/usr/bin/env python
"""bench.py -- D2Q9 MLUPS and achieved HBM GB/s of the fused lattice update on 1/2/4/8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--grid G]
    torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, NCCL)

Workload (BASELINE.json configs[4], the configuration the metric and the north_star targets are
quoted on): the demo_2d_LBM.py double shear layer (BGK, tau from Re = 3000, all cells FLUID,
periodic; demo_2d_LBM.py:26-36, 85-99) on a 32768 x 32768 grid, fp32.  STRONG scaling: the grid
is fixed and split into row slabs over the ranks (jrl_slab.SlabLattice).  A step = one lattice
update of the whole grid = ONE jrl_step_f32 launch per rank (three with slabs: two edge rows +
interior).  rho / u are stored every OUT_EVERY-th step (the demo reads them every 10 steps,
demo_2d_LBM.py:121); the other steps are the pure 72 B per lattice update path.  The demo's
contact-angle wall density only rewrites OBSTACLE cells (LBM_macro_compute_2d.py:271), of which
this grid has none, so it is configured off (identical results).

Prints ONE JSON line (rank 0).  Keys follow the driver contract; see DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200"))

import torch  # noqa: E402

METRIC = "D2Q9 MLUPS (fused stream+collide, fp32) with achieved HBM GB/s vs roofline"
UNIT = "MLUPS"
BYTES_PER_LUP = 72.0     # 9 fp32 reads + 9 fp32 writes; flags / rho / u traffic not counted
OUT_EVERY = 10           # demo_2d_LBM.py:121
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--grid", type=int, default=int(os.environ.get("JRL_BENCH_GRID", 32768)),
                    help="H = W of the periodic shear-layer grid (default: the BASELINE 32768)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end (host-buffer) leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                smax.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# workload: periodic double shear layer (demo_2d_LBM.py:84-99), built slab by slab on the device
# ---------------------------------------------------------------------------------------------
RHO0, VMAX, SIGMA, KAPPA, RE = 0.265, 0.2, 0.05, 80.0, 3000.0


def shear_tau(G):
    return 0.5 + (VMAX * G * 1.0 / RE) / (1.0 / 3.0)   # demo_2d_LBM.py:36


def shear_rows(G, rows, device):
    """rho [1,1,R,G], vel [1,2,R,G] (fp32) of global rows `rows` (a LongTensor, already wrapped)."""
    j = rows.to(torch.float64)
    i = torch.arange(G, dtype=torch.float64)
    ux = torch.where(j <= G / 2.0, VMAX * torch.tanh(KAPPA * (j / G - 0.25)), VMAX * torch.tanh(KAPPA * (-j / G + 0.75)))
    uy = SIGMA * VMAX * torch.sin(2.0 * math.pi * (i / G + 0.25))
    R = rows.numel()
    vel = torch.empty((1, 2, R, G), dtype=torch.float32, device=device)
    vel[0, 0] = ux.to(torch.float32).to(device).view(R, 1)
    vel[0, 1] = uy.to(torch.float32).to(device).view(1, G)
    rho = torch.full((1, 1, R, G), RHO0, dtype=torch.float32, device=device)
    return rho, vel


class Workload:
    """This rank's share of the grid, resident in HBM: populations (double-buffered), flags, and
    the rho / u output buffers."""

    def __init__(self, G, rank, world, device):
        import jrl_slab
        import jrl_step
        from src.LBM.simulation import SimulationParameters, SimulationRunner
        self.G, self.rank, self.world, self.device = G, rank, world, device
        P = SimulationParameters(dim=2, dtype=torch.float32, device=device, simulation_size=(1, 1, G, G), dt=1.0,
                                 density_gas=0.038, density_fluid=0.265, contact_angle=None, Q=9, tau=shear_tau(G), k=0.33)
        self.runner = SimulationRunner(P)
        self.fs = jrl_step.FusedStep(P, operator="bgk", dx=1.0, dt=1.0)
        self.slab = world > 1
        self.row0, self.nrows = jrl_slab.slab_rows(G, world, rank)
        if self.slab:
            rows = torch.arange(self.row0 - 1, self.row0 + self.nrows + 1) % G
        else:
            rows = torch.arange(G)
        rho, vel = shear_rows(G, rows, device)
        col = self.runner.create_collision()
        self.f0_fn = lambda: col.get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel)   # jrl_feq_f32, on the device
        self.flags = torch.ones((1, 1, rows.numel(), G), dtype=torch.uint8, device=device)
        f = self.f0_fn()
        self.rho_out, self.vel_out = rho, vel   # reused as the output buffers of the store steps
        if self.slab:
            self.lat = jrl_slab.SlabLattice(self.fs, G, G, rank=rank, world=world, device=device)
            self.lat.set_state(f, flags=self.flags)
        else:
            self.f, self.f2 = f, torch.empty_like(f)
        self.step_no = 0
        self.cells = self.nrows * G

    def step(self, flags=None, sums=None):
        store = (self.step_no % OUT_EVERY) == OUT_EVERY - 1
        flags = self.flags if flags is None else flags
        if self.slab:
            self.lat.step(flags=flags, sums=sums, store_moments=store, rho_out=self.rho_out, vel_out=self.vel_out) if store else \
                self.lat.step(flags=flags, sums=sums)
        else:
            self.fs.step(self.f, flags, None, None, f_out=self.f2, store_moments=store, sums=sums,
                         rho_out=self.rho_out if store else None, vel_out=self.vel_out if store else None)
            self.f, self.f2 = self.f2, self.f
        self.step_no += 1

    def populations(self):
        return self.lat.owned(self.lat.f) if self.slab else self.f


# ---------------------------------------------------------------------------------------------
# the reference's CPU path on a bounded sample of the same workload (cpu_baseline, --impl reference)
# ---------------------------------------------------------------------------------------------
def cpu_reference_run(G_full, steps, warmup, sample_grid=None, budget_s=20.0):
    """Runs the double shear layer on a `sample_grid`^2 periodic grid (same formula, same tau as the
    full workload) on the host cores and returns (MLUPS, info).  K1 is the reference's own kernel
    body compiled for the CPU when oracle/_ref/libjrl_ref.so exists (kind 'reference'), else the
    numpy oracle (kind 'port')."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    cpu = torch.device("cpu")
    tau = shear_tau(G_full)
    import ref_cpu
    kind = "reference" if ref_cpu.have_ref_lib() else "port"
    Gs = sample_grid or (2048 if kind == "reference" else 512)
    rho, vel = shear_rows(Gs, torch.arange(Gs), cpu)
    flags = torch.ones((1, 1, Gs, Gs), dtype=torch.uint8)
    if kind == "reference":
        import ref_torch_port as RP
        loop = RP.ReferenceCpuLoop(tau)
        cores = loop.cores
        f = RP.get_feq(1.0, 1.0, rho, vel)
        state = [f, rho, vel]

        def one():
            state[0], state[1], state[2] = loop.step(state[0], state[1], state[2], flags)
        what = ("K1 = reference kernel body (propagation.cu) compiled for the host with OpenMP incl. the wrapper's f copy; "
                "moments + feq + BGK = the reference's torch-op chain on CPU tensors")
    else:
        import lbm_oracle as O
        cores = 1
        st = [O.get_feq(1.0, 1.0, rho.numpy(), vel.numpy()), rho.numpy(), vel.numpy()]
        fl = flags.numpy()

        def one():
            fn = O.stream(fl, st[0])
            st[1], st[2] = O.macro_compute(1.0, 1.0, fn, st[1], st[2], fl)
            st[0] = O.collide_bgk(1.0, 1.0, fn, st[1], st[2], fl, tau)
        what = "numpy oracle (scalar port of the reference's algorithm)"
    for _ in range(max(1, warmup)):
        one()
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        one()
        done += 1
        if time.perf_counter() - t0 > budget_s and done >= 2:
            break
    dt = time.perf_counter() - t0
    mlups = Gs * Gs * done / dt / 1e6
    info = {"value": mlups, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{Gs}x{Gs} periodic sub-problem of the {G_full}^2 shear layer (same formula and tau), {done} steps; {what}"}
    return mlups, info, dt / done * 1e3, done


def run_e2e(wl, K, Wm, device, barrier):
    """The same steps through the public API (FusedStep / SlabLattice) with HOST buffers in the loop:
    every step uploads that step's flags field (uint8 [rows, G], authored on the host, pinned memory; a
    host-driven moving obstacle rewrites it every step as the JRL demos do, demo_2d_LBM_JRL_fluid_
    cylinder.py:212) and downloads the step's result, the mass / momentum sums reduced inside the step
    kernel (jrl_step_desc.sums).  Uploads run one step ahead on a copy stream into a double buffer; the
    downloads are asynchronous into a pinned ring and all have landed when the timed region closes.
    Returns (seconds for K steps on this rank, h2d bytes/step, d2h bytes/step, host_sums[K,4])."""
    import jrl_native
    S = jrl_native.SUM_SLOTS
    total = K + Wm
    flags_host = torch.ones(wl.flags.shape, dtype=torch.uint8).pin_memory()
    bufs = [torch.empty_like(wl.flags), torch.empty_like(wl.flags)]
    copy_stream = torch.cuda.Stream(device)
    main = torch.cuda.current_stream(device)
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    sums = torch.zeros((total, S, 4), dtype=torch.float64, device=device)
    host_sums = torch.empty((total, S, 4), dtype=torch.float64).pin_memory()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def upload(i):
        with torch.cuda.stream(copy_stream):
            if i >= 2:
                copy_stream.wait_event(consumed[i % 2])
            bufs[i % 2].copy_(flags_host, non_blocking=True)
            ready[i % 2].record(copy_stream)

    upload(0)
    for i in range(total):
        if i == Wm:
            barrier()
            e0.record()
            upload(i)                    # the timed region holds all K uploads
        if i + 1 < total and i + 1 != Wm:
            upload(i + 1)
        main.wait_event(ready[i % 2])
        wl.step(flags=bufs[i % 2], sums=sums[i])
        consumed[i % 2].record(main)
        host_sums[i].copy_(sums[i], non_blocking=True)
    e1.record()
    barrier()
    return e0.elapsed_time(e1) * 1e-3, flags_host.numel(), S * 4 * 8, host_sums[Wm:].sum(dim=1)


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path (see cpu_reference_run)
    on rank 0; other ranks exit 0 without work."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    mlups, info, ms, done = cpu_reference_run(args.grid, args.steps, args.warmup, budget_s=120.0)
    line = {"impl": "reference", "metric": METRIC, "value": mlups, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"demo_2d_LBM.py double shear layer, D2Q9 BGK, periodic, {args.grid}x{args.grid} fp32 "
                                   f"(reference timed on a bounded sample, see cpu_baseline.sample)", "grid": args.grid},
            "cpu_baseline": info,
            "e2e": {"value": mlups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device. This package has no CPU path (use --impl reference for the CPU reference arm).")
    import torch.distributed as dist
    import jrl_native
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    jrl_native.load()
    G, K, Wm = args.grid, args.steps, max(args.warmup, 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wl = Workload(G, rank, world, device)
    for _ in range(Wm):
        wl.step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = jrl_native.launch_count
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for _ in range(K):
            wl.step()
        e1.record()
        barrier()
    launches = jrl_native.launch_count - launches0
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    nl = torch.tensor([launches], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(nl, op=dist.ReduceOp.SUM)
    total_ms = float(ms.item())
    ms_per_step = total_ms / K
    mlups = G * G * K / (total_ms * 1e-3) / 1e6

    # sanity of the timed state: finite, and total mass conserved (periodic, all FLUID)
    mass = wl.populations().double().sum()
    if world > 1:
        dist.all_reduce(mass)
    mass_rel = abs(float(mass.item()) / (RHO0 * G * G) - 1.0)
    if not (mass_rel < 1e-4):
        raise SystemExit(f"bench.py: mass drifted by {mass_rel:.3e} -- the timed state is wrong")

    e2e = None
    if not args.no_e2e:
        sec, h2d, d2h, hs = run_e2e(wl, K, Wm, device, barrier)
        agg = torch.tensor([sec, float(h2d), float(d2h)], dtype=torch.float64, device=device)
        mass_steps = hs[:, 0].to(device)
        if world > 1:
            t = agg[0:1].clone()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(agg)
            agg[0] = t[0]
            dist.all_reduce(mass_steps)
        worst = float((mass_steps / (RHO0 * G * G) - 1.0).abs().max().item())
        if not (worst < 1e-4):
            raise SystemExit(f"bench.py: e2e per-step mass read back from the device is off by {worst:.3e}")
        e2e = {"value": G * G * K / float(agg[0].item()) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(agg[1].item()),
               "d2h_bytes_per_step": int(agg[2].item()), "ms_per_step": float(agg[0].item()) / K * 1e3,
               "what": "FusedStep/SlabLattice API; per step: H2D of the step's uint8 flags field from pinned host memory "
                       "(prefetched one step ahead on a copy stream), one fused launch, D2H of the in-kernel mass/momentum "
                       "sums (checked on the host: mass conserved to %.1e)" % worst}

    peak, peak_src = measured_peak()
    traffic = None
    try:  # DRAM bytes per lattice update of the dominant kernel from the committed ncu capture, scaled to this launch
        with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as fh:
            traffic = float(json.load(fh)["dram_bytes_per_lup"]) * G * G / world
    except Exception:
        pass
    achieved = BYTES_PER_LUP * G * G / world / (ms_per_step * 1e-3) / 1e9   # per GPU: its cells / the step time
    roofline = {"bound": "hbm", "kernel": "jrl::pipe::k_step_pipe<OP_BGK, 0, 4, 2, 2> (jrl_step_f32; 512-cell tiles, TMA ring, producer warp)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src, "traffic": traffic,
                "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum per lattice update at 8192^2 (profiles/r1_traffic.json) x cells of this launch",
                "algorithmic_bytes_per_launch": BYTES_PER_LUP * G * G / world,
                "note": "per GPU; 72 B per lattice update x cells per rank / CUDA-event step time (1 dominant launch per step)"}
    line = {"metric": METRIC, "value": mlups, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"demo_2d_LBM.py double shear layer, D2Q9 BGK, periodic, {G}x{G} fp32 (BASELINE configs[4] strong-scaling grid)",
                       "grid": G, "tau": shear_tau(G), "rows_per_rank": wl.nrows, "decomposition": "row slabs, ring halo exchange (NCCL)" if world > 1 else "single domain",
                       "moments_stored_every": OUT_EVERY,
                       "flags": "static all-FLUID field; its packed tile mask (jrl_tile_mask_u8) is built once in the value leg and rebuilt every step in the e2e leg (flags re-uploaded)",
                       "l2": "working set (2 x %.1f GB of populations per rank) is far larger than the 126 MB L2" % (36.0 * G * G / world / 1e9)},
            "roofline": roofline, "e2e": e2e, "gpu_launches": int(nl.item()), "mass_drift": mass_rel}
    if rank == 0:
        line["clocks"] = clocks.summary()
        if not args.no_cpu_baseline and world == 1:
            _, info, _, _ = cpu_reference_run(G, steps=8, warmup=1, budget_s=15.0)
            line["cpu_baseline"] = info
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
