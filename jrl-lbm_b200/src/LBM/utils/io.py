"""Host-side file formats of the reference (utils/io.py): Tecplot POINT zones (`.dat`), their
readers, and PNG snapshots.  Off the per-step path; the files are byte
for byte what the reference writes for the same arrays (tests/test_io_formats.py against files the
reference's own writers produced, tests/golden/io/).

Difference in mechanism, not in output: the reference formats one element at a time, which for the
demos' CUDA tensors is one device->host synchronisation per number (4 x H x W per file); here a field
crosses to the host once (or arrives already on the host from `jrl_io.AsyncSnapshotter`) and the rows
are formatted from Python floats.  cv2 is imported lazily: the solver does not need it."""
import os

import numpy as np


def mkdir(path):
    os.makedirs(path, exist_ok=True)


def save_img(tensor_input, filename):
    """Min-max normalised PNG of a [B,C,H,W] tensor, y axis up (utils/io.py:12-29)."""
    if tensor_input.dim() != 4:
        raise RuntimeError("To save an image, the tensor shape should be 4 or 5")
    import cv2
    lo, hi = tensor_input.min(), tensor_input.max()
    img = ((tensor_input - lo) / (hi - lo)).detach().cpu().permute(0, 2, 3, 1) * 255
    cv2.imwrite(filename, cv2.flip(img[0].numpy(), 0))


def _host(a):
    """numpy view of a numpy array or of a (possibly CUDA) torch tensor: one transfer per field."""
    if isinstance(a, np.ndarray):
        return a
    return a.detach().cpu().numpy()


def _centre(np_u, np_v, shape):
    # staggered -> cell centre, in the arrays' own precision as the reference does (utils/io.py:315-318)
    if np_u.shape[-1] == shape[-1] + 1:
        np_u = 0.5 * (np_u[..., 1:] + np_u[..., :-1])
    if np_v.shape[-2] == shape[-2] + 1:
        np_v = 0.5 * (np_v[..., 1:, :] + np_v[..., :-1, :])
    return np_u, np_v


def _column(plane):
    """One [H, W] plane as the strings `"{}".format(x)` gives for its elements: numpy float32 / float64 scalars
    and 0-d torch tensors all format as the Python float of the same value."""
    return [repr(v) for v in np.asarray(plane, dtype=np.float64).ravel().tolist()]


def _write_zone(filename, variables, shape, columns, tail=None):
    H, W = shape[-2], shape[-1]
    xs = [str(i) for i in range(W)] * H
    ys = [s for j in range(H) for s in [str(j)] * W]
    cols = [xs, ys] + columns
    if tail is not None:
        cols += tail
    with open(filename, "w") as fo:
        fo.write('TITLE ="Magnetic internal force"\n')
        fo.write("VARIABLES = {}\n".format(variables))
        fo.write("ZONE I={}, J={}\n".format(W, H))
        fo.write("F=POINT\n")
        fo.write("\n".join(" ".join(r) for r in zip(*cols)))
        if H * W:
            fo.write("\n")


def dump_2d_plt_file_C_rho(filename, np_C, np_density, np_u, np_v, B, C):
    """Tecplot POINT zone: X Y C RHO U V NormX NormY (utils/io.py:308-333)."""
    np_C, np_density, np_u, np_v = _host(np_C), _host(np_density), _host(np_u), _host(np_v)
    shape = np_density.shape
    np_u, np_v = _centre(np_u, np_v, shape)
    H, W = shape[-2], shape[-1]
    nx = [repr(i / 12.5) for i in range(W)] * H
    ny = [s for j in range(H) for s in [repr(j / 12.5)] * W]
    _write_zone(filename, '"X" "Y" "C" "RHO" "U" "V" "NormX" "NormY"', shape,
                [_column(np_C[B, C]), _column(np_density[B, C]), _column(np_u[B, C]), _column(np_v[B, C])], [nx, ny])


def dump_2d_plt_file_single(filename, np_density, np_u, np_v, B, C):
    """Tecplot POINT zone: X Y RHO U V from staggered u, v (utils/io.py:336-352)."""
    np_density, np_u, np_v = _host(np_density), _host(np_u), _host(np_v)
    shape = np_density.shape
    np_u = 0.5 * (np_u[..., 1:] + np_u[..., :-1])
    np_v = 0.5 * (np_v[..., 1:, :] + np_v[..., :-1, :])
    _write_zone(filename, '"X" "Y" "RHO" "U" "V" ', shape,
                [_column(np_density[B, C]), _column(np_u[B, C]), _column(np_v[B, C])])


def dump_2d_plt_file_balance(filename, np_C, np_vel, np_h, B):
    """Tecplot POINT zone: X Y C U V F0..F8 (utils/io.py:222-251).  As in the reference, U and V are
    channels 1 and 2 of `np_vel` (a 3-channel array)."""
    np_C, np_vel, np_h = _host(np_C), _host(np_vel), _host(np_h)
    _write_zone(filename, '"X" "Y" "C" "U" "V" "F0" "F1" "F2" "F3" "F4" "F5" "F6" "F7" "F8"', np_C.shape,
                [_column(np_C[B, 0]), _column(np_vel[B, 1]), _column(np_vel[B, 2])] + [_column(np_h[B, q]) for q in range(9)])


def _read_zone(filename):
    with open(filename, "r") as fo:
        fo.readline()
        fo.readline()
        zone = fo.readline()
        fo.readline()
        i0, j0 = zone.find("I="), zone.find(", J=")
        W, H = int(zone[i0 + 2:j0]), int(zone[j0 + 4:])
        body = np.array([line.split(" ") for line in fo.read().splitlines() if line], dtype=np.float64)
    if body.shape[0] != H * W:
        raise RuntimeError("{}: {} data lines for a {} x {} zone".format(filename, body.shape[0], W, H))
    return H, W, body


def read_2d_plt_file_balance(filename, np_C, np_vel, np_h, B):
    """Inverse of dump_2d_plt_file_balance into caller-provided arrays; U, V land in channels 0, 1
    (utils/io.py:254-278)."""
    H, W, body = _read_zone(filename)
    if (H, W) != tuple(np_C.shape[-2:]):
        raise RuntimeError("{}: zone {} x {} does not match the arrays".format(filename, W, H))
    np_C[B, 0] = body[:, 2].reshape(H, W)
    np_vel[B, 0] = body[:, 3].reshape(H, W)
    np_vel[B, 1] = body[:, 4].reshape(H, W)
    for q in range(9):
        np_h[B, q] = body[:, 5 + q].reshape(H, W)
    return [np_C, np_vel, np_h]


def read_2d_plt_file_C_rho(filename):
    """[C, RHO, U, V] as float32 [1,1,H,W] arrays from a dump_2d_plt_file_C_rho file (utils/io.py:281-305)."""
    H, W, body = _read_zone(filename)
    return [body[:, k].reshape(1, 1, H, W).astype(np.float32) for k in (2, 3, 4, 5)]
