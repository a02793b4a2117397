#!/usr/bin/env python
"""Experiment: a few steps of the small native configs, to be run under `ncu --metrics gpu__time_duration.sum`
for per-launch device times (C1/C2 at 130 x 130, C3 at 800 x 400, fused step via jrl_run_f32)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies
import torch
import jrl_native
jrl_native.load()
import jrl_scenarios as S
api = S.api_from_src()
dev = torch.device("cuda:0")
S.shear_layer(api, dev, res=(130, 130), steps=6, operator="bgk", fused="run")
S.shear_layer(api, dev, res=(130, 130), steps=6, operator="mrt", kbc_type=133, fused="run")
S.cylinder(api, dev, res=(800, 400), D=20, steps=6, fused="run")
torch.cuda.synchronize()
