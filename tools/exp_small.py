#!/usr/bin/env python
"""Experiment: small-grid BGK step time (jrl_run_f32) for builds of the library given on the command line."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))  # tests/jrl_scenarios.py: the demo loop bodies
import torch
import jrl_native
def run(lib):
    jrl_native._lib = None; jrl_native.LIB_PATH = lib; jrl_native.load()
    import jrl_scenarios as S, jrl_step
    api = S.api_from_src(); dev = torch.device("cuda:0")
    res = {}
    for (H, W) in ((130, 130), (800, 400), (512, 512)):
        for kind in ("fluid", "mixed"):
            P = api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.7, contact_angle=None)
            fs = jrl_step.FusedStep(P, operator="bgk")
            st = S.random_state((H, W), 3, dev)
            flags = torch.ones_like(st["flags"]) if kind == "fluid" else st["flags"]
            if kind == "mixed":
                flags = torch.ones_like(st["flags"]); flags[:, :, :, 0] = 2; flags[:, :, :, -1] = 2; flags[:, :, 0, :] = 8; flags[:, :, -1, :] = 16
                flags[:, :, H // 3:H // 3 + 20, W // 3:W // 3 + 20] = 2
            kw = dict(inflow_density=1.0) if kind == "mixed" else {}
            f, r, v = st["f"].clone(), st["rho"].clone(), st["vel"].clone()
            fs.run(f, flags, r, v, 50, force_vec=1, **kw); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fs.run(f, flags, r, v, 400, force_vec=1, **kw); e1.record(); torch.cuda.synchronize()
            res[f"{H}x{W}_{kind}"] = round(e0.elapsed_time(e1) / 400 * 1e3, 2)
    return res
for lib in sys.argv[1:]:
    print(os.path.basename(lib), "us/step:", run(os.path.abspath(lib)), flush=True)
