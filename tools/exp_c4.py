#!/usr/bin/env python
"""Experiment: where does the time of the scaled C4 concentration step (8192^2) go?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies; sys.path.insert(1, os.path.join(ROOT, "tools"))
import torch
import jrl_native
jrl_native.load()
import jrl_scenarios as S
import bench_configs as BC
api = S.api_from_src()
dev = torch.device("cuda:0")
G = int(os.environ.get("G", 8192))
flags, mesh, phi, vel, C = BC.c4_state(api, S, dev, G, G, torch.arange(G), G // 128)
allf = torch.ones_like(flags)
def t(label, axisym=1, fl=flags, ph=phi, clamp=True, inflow=0, fv=0, zero_vel=False):
    P = api.SimulationParameters(simulation_size=(1, 1, G, G), device=dev, contact_angle=None, axisymmetric_type=axisym)
    R = api.SimulationRunner(P)
    fs = R.create_fused_step(operator="bgk", is_convection=True, tau_D=1.0)
    v = torch.zeros_like(vel) if zero_vel else vel
    h = R.create_collision().get_feq_(dx=1.0, dt=1.0, rho=C, vel=torch.zeros_like(vel))
    st = [h, torch.empty_like(h), C.clone(), torch.empty_like(C)]
    def step():
        fs.step(st[0], fl, st[2], v, f_out=st[1], rho_out=st[3], phi_obs=ph, mesh_grid=mesh, inflow_density=inflow, clamp01=clamp, force_vec=fv)
        st[0], st[1], st[2], st[3] = st[1], st[0], st[3], st[2]
    us = BC.fused_loop_us(step, 3, 10)
    print(f"{label:60s} {us:9.1f} us/step  {G*G/us:9.1f} MLUPS", flush=True)
t("as configured (x-axis symmetry, tip block, phi)")
t("register kernel (force_vec=4)", fv=4)
t("pipelined kernel (force_vec=8)", fv=8)
t("pipelined kernel, 2 cells / thread, 3 stages (force_vec=9)", fv=9)
t("pipelined 2c/3s, all FLUID flags, phi", fl=allf, fv=9)
t("pipelined, all FLUID flags, phi", fl=allf, fv=8)
t("pipelined, all FLUID, no phi, no symmetry", fl=allf, ph=None, axisym=0, fv=8)
nt = flags.clone(); nt[flags == 2] = 1; nt[flags == 8] = 1
t("register, geometry without the tip block", fl=nt)
t("pipelined, geometry without the tip block", fl=nt, fv=8)
t("all FLUID flags, phi", fl=allf)
t("all FLUID flags, no phi", fl=allf, ph=None)
t("all FLUID, no phi, no symmetry (CONV 3)", fl=allf, ph=None, axisym=0)
t("all FLUID, no phi, no symmetry, zero velocity", fl=allf, ph=None, axisym=0, zero_vel=True)
t("as configured, zero velocity", zero_vel=True)
