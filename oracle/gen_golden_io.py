#!/usr/bin/env python
"""Writes tests/golden/io/*: the reference's own file writers (src/LBM/utils/io.py:222-352) run on
small seeded arrays, for numpy AND torch inputs (the demos pass torch tensors, whose elements format
as Python doubles; numpy float32 elements format with float32's shortest repr), and the record /
balance lines the demos write (demo_2d_LBM_JRL_C_cylinder.py:245-250, 383-393).

Runs only in the build container (needs /root/reference).  `imageio` is absent from this image and is
imported by the reference module for an unrelated function: it is stubbed.  Test infrastructure only.

    python oracle/gen_golden_io.py"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_IO = "/root/reference/src/LBM/utils/io.py"
OUT = os.path.join(ROOT, "tests", "golden", "io")


def load_reference_io():
    sys.modules.setdefault("imageio", types.ModuleType("imageio"))
    spec = importlib.util.spec_from_file_location("ref_io", REF_IO)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def arrays(seed=11, H=5, W=7):
    """Values that exercise the formatting: tiny / large magnitudes, negative zero, exact integers."""
    r = np.random.RandomState(seed)
    C = r.rand(1, 1, H, W).astype(np.float32)
    C[0, 0, 0, 0] = 0.0
    C[0, 0, 0, 1] = 1.0
    C[0, 0, 0, 2] = 1e-12
    C[0, 0, 0, 3] = -0.0
    C[0, 0, 1, 0] = 123456.789
    rho = (r.rand(1, 1, H, W).astype(np.float32) - 0.5) * 3
    u = (r.rand(1, 1, H, W + 1).astype(np.float32) - 0.5) * 1e-3   # staggered
    v = (r.rand(1, 1, H + 1, W).astype(np.float32) - 0.5) * 1e-3
    uc = (r.rand(1, 1, H, W).astype(np.float32) - 0.5)              # already cell-centred
    vc = (r.rand(1, 1, H, W).astype(np.float32) - 0.5)
    vel3 = r.rand(1, 3, H, W).astype(np.float32)                     # dump_2d_plt_file_balance reads channels 1, 2
    h = r.rand(1, 9, H, W).astype(np.float32)
    return dict(C=C, rho=rho, u=u, v=v, uc=uc, vc=vc, vel3=vel3, h=h)


def main():
    io = load_reference_io()
    os.makedirs(OUT, exist_ok=True)
    a = arrays()
    np.savez(os.path.join(OUT, "inputs.npz"), **a)
    t = {k: torch.from_numpy(x) for k, x in a.items()}
    for kind, d in (("np", a), ("torch", t)):
        io.dump_2d_plt_file_C_rho(os.path.join(OUT, f"c_rho_staggered_{kind}.dat"), d["C"], d["rho"], d["u"], d["v"], 0, 0)
        io.dump_2d_plt_file_C_rho(os.path.join(OUT, f"c_rho_centred_{kind}.dat"), d["C"], d["rho"], d["uc"], d["vc"], 0, 0)
        io.dump_2d_plt_file_single(os.path.join(OUT, f"single_{kind}.dat"), d["rho"], d["u"], d["v"], 0, 0)
        io.dump_2d_plt_file_balance(os.path.join(OUT, f"balance_{kind}.dat"), d["C"], d["vel3"], d["h"], 0)
    # record / balance lines exactly as the demo formats them (torch scalars through .item())
    rec = torch.from_numpy(np.random.RandomState(5).rand(6, 4).astype(np.float32) - 0.3)
    dt_conv_real, vel_obs_real = 1.7e-7, 2.5e-6
    with open(os.path.join(OUT, "record.txt"), "w") as fo:
        fo.write("Step Inflow Current u, v\n")
        for i in range(rec.shape[0]):
            step = 10 * i + 9
            fo.write("{} {} {} {} {}\n".format((step + 1) * dt_conv_real * vel_obs_real * 1e6, rec[i, 0].item(), rec[i, 1].item(),
                                              rec[i, 2].item(), rec[i, 3].item()))
    with open(os.path.join(OUT, "balance.txt"), "w") as fo:
        fo.write("Step Inflow\n")
        for i in range(rec.shape[0]):
            fo.write("{} {}\n".format((10 * i + 1) * dt_conv_real, rec[i, 0].item()))
    np.savez(os.path.join(OUT, "record_inputs.npz"), rec=rec.numpy(), dt_conv_real=dt_conv_real, vel_obs_real=vel_obs_real)
    # the readers, run on the reference's own files
    C0, vel0, h0 = np.zeros_like(a["C"]), np.zeros((1, 2) + a["C"].shape[2:], np.float32), np.zeros_like(a["h"])
    rb = io.read_2d_plt_file_balance(os.path.join(OUT, "balance_np.dat"), C0, vel0, h0, 0)
    rc = io.read_2d_plt_file_C_rho(os.path.join(OUT, "c_rho_centred_np.dat"))
    np.savez(os.path.join(OUT, "read_back.npz"), bal_C=rb[0], bal_vel=rb[1], bal_h=rb[2], C=rc[0], rho=rc[1], u=rc[2], v=rc[3])
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
