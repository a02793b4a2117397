"""Experiment (GPU box): what does the reference's OUTFLOW data race (propagation.cu:195-255) resolve to on a
B200 at the demos' native 128 x 64 grid?  An OUTFLOW cell copies its inward neighbour's f_new while another thread
streams into it.  For every OUTFLOW cell of the stage-A flags layout the reference kernel's result (prebuilt extension
from baseline/_ref) is classified over repeated launches as `post` (= the neighbour's value after streaming: this
library's definition), `pre` (= the neighbour's previous-step populations, the wrapper's f + 0.0 copy) or `mixed`."""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200"))
so = os.path.join(ROOT, "baseline", "_ref", "ext", "propagation", "propagation.so")
spec = importlib.util.spec_from_file_location("propagation", so)
prop_ext = importlib.util.module_from_spec(spec)
spec.loader.exec_module(prop_ext)
import jrl_lazy  # noqa: E402
jrl_lazy.ENABLED = False
from src.LBM.simulation import SimulationParameters, SimulationRunner  # noqa: E402

dev = torch.device("cuda:0")
out = {}
for (H, W), label in (((128, 64), "stage A 128x64"), ((800, 400), "C3 800x400 layout (OUTFLOW top row)")):
    flags = torch.full((1, 1, H, W), 16, dtype=torch.uint8, device=dev)
    flags[..., 1:-1, 1:-1] = 1
    if H == 128:
        flags[..., 0] = 1
        flags[..., -1, :] = 2
        flags[..., 0:21, 0:26] = 2
    else:
        flags[...] = 2
        flags[..., 1:-1, 1:-1] = 1
        flags[..., 0, 1:-1] = 8
        flags[..., -1, 1:-1] = 16
    g = torch.Generator().manual_seed(7)
    f = (torch.rand((1, 9, H, W), generator=g) * 0.2 + 0.05).to(dev)
    rho = torch.ones((1, 1, H, W), device=dev)
    vel = torch.zeros((1, 2, H, W), device=dev)
    P = SimulationParameters(dim=2, dtype=torch.float32, device=dev, simulation_size=(1, 1, H, W), dt=1.0, contact_angle=None, Q=9,
                             tau=1.0, axisymmetric_type=1 if H == 128 else 0)
    ours = SimulationRunner(P).create_propagation().propagation(f=f, flags=flags, rho=rho.clone(), vel=vel.clone())
    ours = ours[2] if isinstance(ours, tuple) else ours
    ys, xs = np.where(flags[0, 0].cpu().numpy() == 16)
    res = {"cells": int(len(ys)), "runs": []}
    for run in range(10):
        f_new = f + 0.0
        prop_ext.propagation_2d_forward(flags, f, torch.Tensor([]).to(dev), f_new, rho.clone(), vel.clone(), torch.Tensor([]).to(dev),
                                        -1.0, False, 1 if H == 128 else 0)
        torch.cuda.synchronize()
        r, o, p = f_new.cpu().numpy()[0], ours.cpu().numpy()[0], f.cpu().numpy()[0]
        cnt = {}
        for y, x in zip(ys, xs):
            sy = y + 1 if y == 0 else (y - 1 if y == H - 1 else y)
            sx = x + 1 if x == 0 else (x - 1 if x == W - 1 else x)
            if (sy, sx) == (y, x):
                continue
            where = "row" if sy != y and sx == x else ("col" if sx != x and sy == y else "corner")
            npost, npre = int((r[:, y, x] == o[:, y, x]).sum()), int((r[:, y, x] == p[:, sy, sx]).sum())
            k = "post" if npost == 9 else ("pre" if npre == 9 else f"mixed")
            cnt[f"{where}:{k}"] = cnt.get(f"{where}:{k}", 0) + 1
        res["runs"].append(cnt)
    out[label] = res
print(json.dumps(out, indent=1))
with open(os.path.join(ROOT, "gpurun_out", "outflow_race.json"), "w") as fh:
    json.dump(out, fh, indent=1)
