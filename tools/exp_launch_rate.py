#!/usr/bin/env python
"""Experiment: is the 130 x 130 run loop bound by the host's enqueue rate or by the device?  Times the jrl_run_f32 call
on the host (returns when everything is enqueued) and the same steps on the device (CUDA events)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.insert(0, ROOT)
import torch
import jrl_native
jrl_native.load()
import jrl_step
from src.LBM.simulation import SimulationParameters
from jrl_solver import LBMCollision2d
dev = torch.device("cuda:0")
for (H, W) in ((130, 130), (128, 64), (512, 512)):
    for op in ("bgk", "kbc"):
        P = SimulationParameters(device=dev, simulation_size=(1, 1, H, W), contact_angle=None, tau=0.6)
        fs = jrl_step.FusedStep(P, operator=op, kbc_type=133 if op == "kbc" else None)
        rho = torch.ones((1, 1, H, W), device=dev); vel = torch.zeros((1, 2, H, W), device=dev)
        vel[:, 0] = 0.05 * torch.rand((H, W), device=dev)
        f = LBMCollision2d(device=dev).get_feq_(1.0, 1.0, rho, vel)
        f2 = torch.empty_like(f)
        flags = torch.ones((1, 1, H, W), dtype=torch.uint8, device=dev)
        N = 20000
        fs.run(f, flags, None, None, 200, f_alt=f2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        fs.run(f, flags, None, None, N, f_alt=f2)
        e1.record(); t1 = time.perf_counter()
        torch.cuda.synchronize(); t2 = time.perf_counter()
        print(json.dumps({"grid": [H, W], "op": op, "host_enqueue_us_per_step": round((t1 - t0) / N * 1e6, 2),
                          "device_us_per_step": round(e0.elapsed_time(e1) / N * 1e3, 2), "wall_us_per_step": round((t2 - t0) / N * 1e6, 2)}), flush=True)
