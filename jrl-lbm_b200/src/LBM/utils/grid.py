"""Grid helpers the demos use to build per-step inputs (utils/grid.py:7-82 of the reference).
Only what the 2-D hot path's drivers call is provided (the 3-D / multiphase
helpers are unused)."""
import torch
import torch.nn.functional as F


def create_2d_meshgrid_tensor(size, device=torch.device("cpu"), dtype=torch.float32):
    """[B,2,H,W]: channel 0 = x index, channel 1 = y index (utils/grid.py:67-82)."""
    batch, _, height, width = size
    ys = torch.arange(0, height, device=device, dtype=dtype).view(height, 1).expand(height, width)
    xs = torch.arange(0, width, device=device, dtype=dtype).view(1, width).expand(height, width)
    return torch.stack([xs, ys], dim=0).unsqueeze(0).repeat(batch, 1, 1, 1)


def _staggered(t, axis, mode):
    if t.dim() != 4:
        raise RuntimeError("A grid has to be 2D [B, C, H, W] to be staggered")
    n = t.shape[axis]
    mid = (t.narrow(axis, 1, n - 1) + t.narrow(axis, 0, n - 1)) * 0.5
    pad = (1, 1, 0, 0) if axis == 3 else (0, 0, 1, 1)
    return F.pad(mid, pad=pad, mode=mode, value=0)


def get_staggered_x(input, mode="replicate"):
    return _staggered(input, 3, mode)


def get_staggered_y(input, mode="replicate"):
    return _staggered(input, 2, mode)


def get_staggered(input, mode="replicate"):
    """Face-centred [u, v] from a cell-centred [B,2,H,W] velocity (utils/grid.py:52-64)."""
    if input.shape[1] != 2:
        raise RuntimeError("Only 2D scene supported")
    return [get_staggered_x(input[:, 0:1, ...], mode), get_staggered_y(input[:, 1:2, ...], mode)]
