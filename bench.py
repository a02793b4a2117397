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
PRE_ROLL_S = 0.6          # untimed steps under load (and under the clock sampler) before the timed region
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
    ap.add_argument("--no-configs", action="store_true", help="skip the extra BASELINE configs (c2 / c3 / c4 at 8192^2)")
    ap.add_argument("--e2e-full-flags", action="store_true", help="e2e leg uploads the whole flags field every step (round-1 behaviour)")
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
    """SM clock and throttle reasons sampled every ~5 ms through NVML (nvidia_ml_py) on a thread, so that even the
    30 ms timed region of the 8-GPU run holds samples; falls back to `nvidia-smi -lms` when NVML is not importable.
    `mark()` brackets the timed region: only samples between the marks are summarised (the region is preceded by at
    least PRE_ROLL_S of untimed steps under the same sampler, so the clocks are the ones under load)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.marks, self.stop = index, [], None, [], False
        self.nvml = None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return self
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _poll(self):
        n = self.nvml
        bits = {"hw_slowdown": n.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": n.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": n.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": n.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop:
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((time.perf_counter(), sm, self.smax, [k for k, b in bits.items() if r & b]))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            c = [x.strip() for x in line.split(",")]
            try:
                self.rows.append((time.perf_counter(), float(c[1]), float(c[2]),
                                  [n for n, v in zip(names, c[5:9]) if v.lower().startswith("active")]))
            except Exception:
                pass

    def mark(self):
        self.marks.append(time.perf_counter())

    def __exit__(self, *a):
        self.stop = True
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
        if getattr(self, "t", None) is not None:
            self.t.join(timeout=2)

    def summary(self):
        rows = self.rows
        inside = [r for r in rows if len(self.marks) >= 2 and self.marks[0] <= r[0] <= self.marks[-1]]
        use = inside if inside else rows
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(r[1] for r in use)
        reasons = sorted({x for r in use for x in r[3]})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(r[2] for r in use), "reasons": reasons,
                "samples": len(inside), "samples_under_load_window": len(rows),
                "source": "nvml, 5 ms period" if self.nvml is not None else "nvidia-smi -lms 20"}


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
    # all host cores this process may use, at every N: torch.distributed.run exports OMP_NUM_THREADS=1, which made the
    # round-1 CPU arm 7x slower under torchrun than alone
    try:
        threads = len(os.sched_getaffinity(0))
    except AttributeError:
        threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    Gs = sample_grid or (2048 if kind == "reference" else 512)
    rho, vel = shear_rows(Gs, torch.arange(Gs), cpu)
    flags = torch.ones((1, 1, Gs, Gs), dtype=torch.uint8)
    if kind == "reference":
        import ref_torch_port as RP
        loop = RP.ReferenceCpuLoop(tau, threads=threads)
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


E2E_BAND_ROWS = 64


def run_e2e(wl, K, Wm, device, barrier, full_flags=False):
    """The same steps through the public API (FusedStep / SlabLattice) with HOST buffers in the loop.  The host owns
    the flags field and a host-driven moving obstacle rewrites part of it every step (as the JRL demos do,
    demo_2d_LBM_JRL_fluid_cylinder.py:205-214): per step the DIRTY ROWS -- a band of E2E_BAND_ROWS rows that moves one
    row per step -- are uploaded from pinned host memory into the device flags field (full_flags=True: the whole uint8
    field, the round-1 behaviour, PCIe-bound at 1 B per cell), the packed tile mask is rebuilt because flags changed,
    the step runs, and its result -- the mass / momentum sums reduced inside the step kernel (jrl_step_desc.sums) -- is
    downloaded into a pinned ring; all downloads have landed when the timed region closes.
    Returns (seconds for K steps on this rank, h2d bytes/step, d2h bytes/step, host_sums[K,4])."""
    import jrl_native
    S = jrl_native.SUM_SLOTS
    total = K + Wm
    rows_dev = wl.flags.shape[2]
    R = rows_dev if full_flags else min(E2E_BAND_ROWS, max(1, rows_dev - 2))
    band_host = torch.ones((1, 1, R, wl.flags.shape[3]), dtype=torch.uint8).pin_memory()
    copy_stream = torch.cuda.Stream(device)
    main = torch.cuda.current_stream(device)
    ready, consumed = torch.cuda.Event(), torch.cuda.Event()
    sums = torch.zeros((total, S, 4), dtype=torch.float64, device=device)
    host_sums = torch.empty((total, S, 4), dtype=torch.float64).pin_memory()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    span = max(1, rows_dev - R - 1)

    def upload(i):
        y0 = 0 if full_flags else 1 + (i % span)
        with torch.cuda.stream(copy_stream):
            if i > 0:
                copy_stream.wait_event(consumed)     # the previous step has read the flags this copy overwrites
            wl.flags[:, :, y0:y0 + R].copy_(band_host, non_blocking=True)
            ready.record(copy_stream)

    for i in range(total):
        if i == Wm:
            barrier()
            e0.record()
        upload(i)
        main.wait_event(ready)
        wl.step(flags=wl.flags, sums=sums[i])
        consumed.record(main)
        host_sums[i].copy_(sums[i], non_blocking=True)
    e1.record()
    barrier()
    return e0.elapsed_time(e1) * 1e-3, band_host.numel(), S * 4 * 8, host_sums[Wm:].sum(dim=1)


def _event_ms(step, K, warm, barrier, world, device):
    import torch.distributed as dist
    for _ in range(warm):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / K], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def extra_configs(device, rank, world, barrier, peak, G2=8192, K=20):
    """The other BASELINE configs at their bandwidth-bound size, measured in the same process
    after the headline: 20 CUDA-event-timed steps each after 3 warm-up steps, inputs (>= 4.8 GB) far larger than L2.
    At N > 1 only c4_conc_scaled (the config BASELINE shards over row slabs) runs.  `frac` = algorithmic bytes per
    step / time / measured HBM peak, with the per-cell byte count stated per entry."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import jrl_slab
    import jrl_step
    import workloads as WL
    from src.LBM.simulation import SimulationParameters, SimulationRunner
    out = {}

    def entry(ms, cells_per_gpu, bytes_per_cell, what, **kw):
        gbs = bytes_per_cell * cells_per_gpu / (ms * 1e-3) / 1e9
        return dict(ms_per_step=ms, mlups=cells_per_gpu * world / ms / 1e3, bytes_per_cell=bytes_per_cell,
                    algorithmic_bytes_per_step=bytes_per_cell * cells_per_gpu, achieved_gbs=gbs, frac=gbs / peak, what=what, **kw)

    if world == 1:
        G = G2
        rho, vel = shear_rows(G, torch.arange(G), device)
        flags = torch.ones((1, 1, G, G), dtype=torch.uint8, device=device)
        for name, op, kt in (("c2_mrt", "mrt", None), ("c2_kbc_a", "kbc", 133)):
            P = SimulationParameters(simulation_size=(1, 1, G, G), device=device, contact_angle=None, tau=shear_tau(G))
            R = SimulationRunner(P)
            fs = R.create_fused_step(operator=op, kbc_type=kt)
            st = [R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel), None]
            st[1] = torch.empty_like(st[0])

            def step():
                fs.step(st[0], flags, None, None, f_out=st[1], store_moments=False)
                st[0], st[1] = st[1], st[0]
            out[name] = entry(_event_ms(step, K, 3, barrier, world, device), G * G, 72.0,
                              f"demo_2d_LBM_KBC.py shear layer at {G}^2, {op.upper()}{'' if kt is None else ' type %d' % kt}, fused step, moments not stored")
            del st
        # the reference's API, unchanged demo loop body (demo_2d_LBM.py:101-119): three solver-object calls per step,
        # which jrl_lazy turns into one fused launch; next to the explicit FusedStep call on the same buffers
        P = SimulationParameters(dim=2, dtype=torch.float32, device=device, simulation_size=(1, 1, G, G), dt=1.0,
                                 contact_angle=None, Q=9, tau=shear_tau(G))
        R = SimulationRunner(P)
        prop, macro, col = R.create_propagation(), R.create_macro_compute(), R.create_collision()
        st = dict(f=col.get_feq_(dx=1.0, dt=1.0, rho=rho, vel=vel), rho=rho.clone(), vel=vel.clone())
        force = torch.zeros((1, 2, G, G), dtype=torch.float32, device=device)

        def demo_step():
            f = prop.propagation(f=st["f"], flags=flags)
            st["rho"], st["vel"] = macro.macro_compute(dx=1.0, dt=1.0, f=f, rho=st["rho"], vel=st["vel"], flags=flags)
            st["f"] = col.collision(dx=1.0, dt=1.0, f=f, rho=st["rho"], vel=st["vel"], flags=flags, force=force, KBC_type=None)
        import jrl_native
        n0 = jrl_native.launch_count
        ms_api = _event_ms(demo_step, K, 3, barrier, world, device)
        per_step = (jrl_native.launch_count - n0) / (K + 3)
        fs = R.create_fused_step(operator="bgk")
        bufs = [st["f"], torch.empty_like(st["f"]), st["rho"], torch.empty_like(st["rho"]), st["vel"], torch.empty_like(st["vel"])]

        def fused_step():
            fs.step(bufs[0], flags, bufs[2], bufs[4], f_out=bufs[1], rho_out=bufs[3], vel_out=bufs[5], force=force)
            bufs[0], bufs[1], bufs[2], bufs[3], bufs[4], bufs[5] = bufs[1], bufs[0], bufs[3], bufs[2], bufs[5], bufs[4]
        ms_fs = _event_ms(fused_step, K, 3, barrier, world, device)
        out["c1_reference_api"] = entry(ms_api, G * G, 84.0, f"demo_2d_LBM.py loop body UNCHANGED (prop.propagation -> macro.macro_compute -> "
                                        f"collision.collision) at {G}^2: {per_step:.2f} library launches per step; 72 B populations + 12 B "
                                        "rho, u stored every step as that API returns them", launches_per_step=per_step,
                                        fused_step_same_buffers_ms=ms_fs, ratio_to_fused_step=ms_api / ms_fs)
        del st, bufs, force
        torch.cuda.empty_cache()
        # C3 scaled: flow past a cylinder, MRT + moving-wall step + inflow / outflow + contact-angle ring
        H, W, D = G, G, 205
        cflags, phi, crho, cvel, inflow_vel = WL.cylinder_state(device, H, W, D, (2048.0, 2048.0))
        tau = WL.cylinder_tau(H, W)
        P = SimulationParameters(dim=2, dtype=torch.float32, device=device, simulation_size=(1, 1, H, W), dt=1.0, density_gas=0,
                                 density_fluid=1.0, contact_angle=torch.Tensor([0.5 * math.pi]).to(device), Q=9, tau=tau)
        R = SimulationRunner(P)
        col = R.create_collision_MRT()
        col.preset_KBC(dx=1.0, dt=1.0, tau=tau)
        fs = R.create_fused_step(operator="mrt")
        f0 = col.get_feq_(dx=1.0, dt=1.0, rho=crho, vel=cvel)
        b = [f0, torch.empty_like(f0), crho, torch.empty_like(crho), cvel, torch.empty_like(cvel)]
        zero = torch.zeros_like(cvel)

        def cyl_step():
            fs.step(b[0], cflags, b[2], b[4], f_out=b[1], rho_out=b[3], vel_out=b[5], inflow_density=1.0, inflow_vel=inflow_vel,
                    use_wall=True, wall_phi=phi, obs_vel=zero)
            b[0], b[1], b[2], b[3], b[4], b[5] = b[1], b[0], b[3], b[2], b[5], b[4]
        out["c3_cylinder_8192"] = entry(_event_ms(cyl_step, K, 3, barrier, world, device), H * W, 72.0,
                                        f"demo_2d_LBM_cylinder.py scaled to {H}x{W} (D = {D}): MRT, moving-wall step, INFLOW / OUTFLOW rows, "
                                        "contact-angle ring, rho / u read and stored every step (+24 B per cell not counted)")
        del b, f0, zero, cflags, phi, crho, cvel
        torch.cuda.empty_cache()
    # C4 scaled: concentration step on the scaled tip geometry, row slabs at N > 1
    G = G2
    P = SimulationParameters(simulation_size=(1, 1, G, G), device=device, contact_angle=None, axisymmetric_type=1)
    R = SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    row0, nrows = jrl_slab.slab_rows(G, world, rank)
    rows = torch.arange(row0 - 1, row0 + nrows + 1) % G if world > 1 else torch.arange(G)
    flags, mesh, phi, vel, C = WL.c4_state(device, G, G, rows, G // 128)
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    kw = dict(phi_obs=phi, mesh_grid=mesh, inflow_density=0, clamp01=True)
    if world > 1:
        # graph=True: the slab step (both streams, NCCL p2p included) is replayed from a CUDA graph -- at 1024 rows per
        # rank the step's ~15 host calls take longer than its kernels (profiles/r2_slab_trace.md); released before the
        # process group is destroyed (jrl_slab.release_all_graphs at the end of main)
        lat = jrl_slab.SlabLattice(fs, G, G, rank=rank, world=world, device=device, graph=os.environ.get("JRL_SLAB_GRAPH", "1") == "1")
        lat.set_state(h, flags=flags, rho=C, vel=vel, **kw)
        Cout = torch.empty_like(C)

        def c4_step():
            lat.step(store_moments=True, rho_out=Cout)
        for _ in range(6):   # both ping-pong parities seen twice: the graphs are captured before the timed steps
            c4_step()
    else:
        st = [h, torch.empty_like(h), C, torch.empty_like(C)]

        def c4_step():
            fs.step(st[0], flags, st[2], vel, f_out=st[1], rho_out=st[3], **kw)
            st[0], st[1], st[2], st[3] = st[1], st[0], st[3], st[2]
    out["c4_conc_scaled"] = entry(_event_ms(c4_step, K, 3, barrier, world, device), nrows * G, 84.0,
                                  f"demo_2d_LBM_JRL_C_cylinder.py scaled to {G}^2 (tip radius 6.5 x {G // 128}, x-axis symmetry, is_negative layout): "
                                  "fused concentration step, static tip, analytic stand-in velocity; 72 B h + 8 B u read + 4 B C write",
                                  decomposition="row slabs, ring halo exchange (NCCL)" if world > 1 else "single domain", n_gpus=world)
    return out


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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for _ in range(max(Wm, 1)):
            wl.step()
        e1.record()
        barrier()
        # untimed pre-roll: PRE_ROLL_S of the same steps, so that the timed region starts with the GPU at its
        # under-load clocks and the sampler has a load window around it (same count on every rank: slabs exchange halos)
        pre = torch.tensor([PRE_ROLL_S * 1e3 / max(e0.elapsed_time(e1) / max(Wm, 1), 1e-3)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(pre, op=dist.ReduceOp.MAX)
        pre = int(min(5000, max(0, math.ceil(float(pre.item())) - max(Wm, 1))))
        for i in range(pre):
            wl.step()
            if i % 32 == 31:
                torch.cuda.synchronize()
        barrier()
        launches0 = jrl_native.launch_count
        clocks.mark()
        e0.record()
        for _ in range(K):
            wl.step()
        e1.record()
        barrier()
        clocks.mark()
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
        sec, h2d, d2h, hs = run_e2e(wl, K, Wm, device, barrier, full_flags=args.e2e_full_flags)
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
               "what": ("FusedStep/SlabLattice API; per step: H2D of the step's dirty flag rows (a %d-row band of the host-owned uint8 "
                        "flags field that moves one row per step, pinned memory -> device flags)" % E2E_BAND_ROWS if not args.e2e_full_flags
                        else "FusedStep/SlabLattice API; per step: H2D of the whole uint8 flags field from pinned host memory")
                       + ", tile-mask rebuild, one fused launch, D2H of the in-kernel mass/momentum sums (checked on the host: "
                         "mass conserved to %.1e)" % worst}

    peak, peak_src = measured_peak()
    traffic = None
    try:  # DRAM bytes per lattice update of the dominant kernel from the committed ncu capture, scaled to this launch
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as fh:
            traffic = float(json.load(fh)["dram_bytes_per_lup"]) * G * G / world
    except Exception:
        pass
    achieved = BYTES_PER_LUP * G * G / world / (ms_per_step * 1e-3) / 1e9   # per GPU: its cells / the step time
    roofline = {"bound": "hbm", "kernel": "jrl::pipe::k_step_pipe<OP_BGK, 0, 4, 2, 2> (jrl_step_f32; 512-cell tiles, TMA ring, producer warp)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src, "traffic": traffic,
                "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum per lattice update at 8192^2 (profiles/r2_traffic.json) x cells of this launch",
                "algorithmic_bytes_per_launch": BYTES_PER_LUP * G * G / world,
                "note": "per GPU; 72 B per lattice update x cells per rank / CUDA-event step time (1 dominant launch per step)"}
    line = {"metric": METRIC, "value": mlups, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"demo_2d_LBM.py double shear layer, D2Q9 BGK, periodic, {G}x{G} fp32 (BASELINE configs[4] strong-scaling grid)",
                       "grid": G, "tau": shear_tau(G), "rows_per_rank": wl.nrows, "decomposition": "row slabs, ring halo exchange (NCCL)" if world > 1 else "single domain",
                       "moments_stored_every": OUT_EVERY,
                       "flags": "static all-FLUID field; its packed tile mask (jrl_tile_mask_u8) is built once in the value leg and rebuilt every step in the e2e leg (dirty rows re-uploaded)",
                       "l2": "working set (2 x %.1f GB of populations per rank) is far larger than the 126 MB L2" % (36.0 * G * G / world / 1e9)},
            "roofline": roofline, "e2e": e2e, "gpu_launches": int(nl.item()), "mass_drift": mass_rel}
    line["config"]["pre_roll_steps"] = pre
    if not args.no_configs:
        del wl
        torch.cuda.empty_cache()
        line["configs"] = extra_configs(device, rank, world, barrier, peak, G2=min(8192, G))
    if rank == 0:
        line["clocks"] = clocks.summary()
        if not args.no_cpu_baseline and world == 1:
            _, info, _, _ = cpu_reference_run(G, steps=8, warmup=1, budget_s=15.0)
            line["cpu_baseline"] = info
        print(json.dumps(line), flush=True)
    if world > 1:
        import jrl_slab
        jrl_slab.release_all_graphs()   # captured NCCL operations must be gone before the communicator is
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
