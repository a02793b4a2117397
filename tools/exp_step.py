"""Experiment harness (not part of the product): time jrl_step_f32 variants on one GPU.
usage: python tools/exp_step.py LIB [LIB ...]   -- each LIB a build of libjrl_lbm.so"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.insert(0, ROOT)
import torch
import jrl_native

def run(lib, G=int(os.environ.get("G", 8192)), steps=20, op="bgk"):
    jrl_native._lib = None
    jrl_native.LIB_PATH = lib
    jrl_native.load()
    import bench, jrl_step
    from src.LBM.simulation import SimulationParameters
    dev = torch.device("cuda:0")
    P = SimulationParameters(device=dev, simulation_size=(1, 1, G, G), contact_angle=None, tau=bench.shear_tau(G))
    fs = jrl_step.FusedStep(P, operator=op, kbc_type=133 if op == "kbc" else None)
    rho, vel = bench.shear_rows(G, torch.arange(G), dev)
    from jrl_solver import LBMCollision2d
    f = LBMCollision2d(device=dev).get_feq_(1.0, 1.0, rho, vel)
    f2 = torch.empty_like(f)
    flags = torch.ones((1, 1, G, G), dtype=torch.uint8, device=dev)
    out = {}
    for vec in [int(v) for v in os.environ.get("VECS", "8,4,2").split(",")]:
        for _ in range(3):
            fs.step(f, flags, None, None, f_out=f2, store_moments=False, force_vec=vec); f, f2 = f2, f
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fs.step(f, flags, None, None, f_out=f2, store_moments=False, force_vec=vec); f, f2 = f2, f
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out[f"vec{vec}"] = round(72.0 * G * G / ms / 1e6, 1)
    return out

if __name__ == "__main__":
    ops = os.environ.get("OPS", "bgk").split(",")
    for lib in sys.argv[1:]:
        for op in ops:
            try:
                print(os.path.basename(lib), op, "GB/s:", json.dumps(run(os.path.abspath(lib), op=op)), flush=True)
            except Exception as e:
                print(os.path.basename(lib), op, "FAILED:", str(e)[:120], flush=True)
