#!/usr/bin/env python
"""Experiment: where does the time of the C3 (cylinder, 800 x 400) fused step go?"""
import os, sys, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies
import torch
import jrl_native
jrl_native.load()
import jrl_scenarios as S
api = S.api_from_src()
dev = torch.device("cuda:0")
res = (int(os.environ.get("H", 800)), int(os.environ.get("W", 400)))
out = S.cylinder(api, dev, res=res, D=20, steps=50, fused="run")
flags = torch.from_numpy(out["flags"]).to(dev)
phi = torch.from_numpy(out["phi_obs"]).to(dev)
f = torch.from_numpy(out["f"]).to(dev); rho = torch.from_numpy(out["rho"]).to(dev); vel = torch.from_numpy(out["vel"]).to(dev)
P = api.SimulationParameters(simulation_size=(1, 1, *res), device=dev, tau=float(out["tau"]), contact_angle=None)
Pc = api.SimulationParameters(simulation_size=(1, 1, *res), device=dev, tau=float(out["tau"]), contact_angle=torch.Tensor([0.5 * math.pi]).to(dev))
R = api.SimulationRunner(P)
iv = torch.tensor([0.0, 0.05], device=dev)
ov = torch.zeros_like(vel)
def t(label, fs, n=200, **kw):
    fs.run(f.clone(), flags, rho.clone(), vel.clone(), 20, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ff, rr, vv = f.clone(), rho.clone(), vel.clone()
    e0.record(); fs.run(ff, flags, rr, vv, n, **kw); e1.record(); torch.cuda.synchronize()
    print(f"{label:50s} {e0.elapsed_time(e1) / n * 1e3:8.2f} us/step", flush=True)
base = dict(inflow_density=1.0, inflow_vel=iv)
wall = dict(use_wall=True, wall_phi=phi, obs_vel=ov)
for vec in (0, 4, 2, 1):
    t(f"mrt + wall, force_vec={vec}", R.create_fused_step(operator="mrt"), force_vec=vec, **base, **wall)
t("mrt + wall + contact ring", api.SimulationRunner(Pc).create_fused_step(operator="mrt"), **base, **wall)
t("mrt, no wall", R.create_fused_step(operator="mrt"), **base)
t("bgk + wall", R.create_fused_step(operator="bgk"), **base, **wall)
t("bgk, no wall", R.create_fused_step(operator="bgk"), **base)
allf = torch.ones_like(flags)
t("mrt, all FLUID flags", R.create_fused_step(operator="mrt"), **{})
flags = allf
t("mrt, all FLUID flags (really)", R.create_fused_step(operator="mrt"))
t("bgk, all FLUID flags", R.create_fused_step(operator="bgk"))
t("bgk, all FLUID, moments as outputs", R.create_fused_step(operator="bgk"))
