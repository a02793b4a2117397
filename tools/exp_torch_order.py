"""Experiment (GPU box): dump inputs and outputs of the torch-CUDA batched 9x9 mat-vec products the reference's MRT
and KBC operators use (LBM_collision_MRT_2d.py:97-100, LBM_collision_KBC_2d.py:100-130), so that the summation order
cuBLAS uses for them can be identified offline (gpurun_out/torch_order.npz)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200"))
from jrl_solver import _MRT_M  # noqa: E402

dev = torch.device("cuda:0")
out = {}
for tau in (0.526, 0.8):
    M = torch.Tensor(_MRT_M).to(torch.float32).to(dev).reshape(1, 1, 1, 9, 9)
    s7 = 1.0 / tau
    s4 = 8.0 * (2 - s7) / (8 - s7)
    S = torch.Tensor([0.0, s7, s7, 0.0, s4, 0.0, s4, s7, s7]).to(torch.float32).to(dev).reshape(1, 1, 1, 9, 1)
    Minv = torch.linalg.inv(M)
    A = Minv @ (S * M)
    out[f"Minv_{tau}"] = Minv.cpu().numpy()
    out[f"A_{tau}"] = A.cpu().numpy()
    out[f"SM_{tau}"] = (S * M).cpu().numpy()
    g = torch.Generator().manual_seed(5)
    for shape in ((1, 9, 24, 20), (1, 9, 130, 130), (2, 9, 16, 12)):
        d = ((torch.rand(shape, generator=g) - 0.5) * 0.1).to(dev)
        om = (A @ d.permute(0, 2, 3, 1).unsqueeze(-1)).squeeze(-1).permute(0, 3, 1, 2)
        key = f"{tau}_{shape[2]}x{shape[3]}x{shape[0]}"
        out["d_" + key] = d.cpu().numpy()
        out["om_" + key] = om.contiguous().cpu().numpy()
# per-cell matrices (KBC A/B central moments): [B,H,W,9,9] @ [B,H,W,9,1]
g = torch.Generator().manual_seed(6)
E = (torch.rand((1, 24, 20, 9, 9), generator=g) - 0.5).to(dev)
x = torch.rand((1, 9, 24, 20), generator=g).to(dev)
y = (E @ x.permute(0, 2, 3, 1).unsqueeze(-1)).squeeze(-1).permute(0, 3, 1, 2)
out["E"], out["x"], out["y"] = E.cpu().numpy(), x.cpu().numpy(), y.contiguous().cpu().numpy()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez(os.path.join(ROOT, "gpurun_out", "torch_order.npz"), **out)
print("wrote gpurun_out/torch_order.npz", torch.__version__, torch.backends.cuda.matmul.allow_tf32)
