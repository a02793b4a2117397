"""Host-side file writers the demo scripts import (utils/io.py of the reference).  Off the
per-step path; same file formats.  cv2 is imported lazily: the solver does
not need it."""
import os


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


def _centre(np_u, np_v, shape):
    if np_u.shape[-1] == shape[-1] + 1:
        np_u = 0.5 * (np_u[..., 1:] + np_u[..., :-1])
    if np_v.shape[-2] == shape[-2] + 1:
        np_v = 0.5 * (np_v[..., 1:, :] + np_v[..., :-1, :])
    return np_u, np_v


def _write_zone(filename, variables, shape, rows):
    with open(filename, "w") as fo:
        fo.write('TITLE ="Magnetic internal force"\n')
        fo.write("VARIABLES = {}\n".format(variables))
        fo.write("ZONE I={}, J={}\n".format(shape[-1], shape[-2]))
        fo.write("F=POINT\n")
        for row in rows:
            fo.write(" ".join(str(v) for v in row) + "\n")


def dump_2d_plt_file_C_rho(filename, np_C, np_density, np_u, np_v, B, C):
    """Tecplot POINT zone: X Y C RHO U V NormX NormY (utils/io.py:308-333)."""
    shape = np_density.shape
    np_u, np_v = _centre(np_u, np_v, shape)
    _write_zone(filename, '"X" "Y" "C" "RHO" "U" "V" "NormX" "NormY"', shape,
                ((i, j, np_C[B, C, j, i], np_density[B, C, j, i], np_u[B, C, j, i], np_v[B, C, j, i], i / 12.5, j / 12.5)
                 for j in range(shape[-2]) for i in range(shape[-1])))


def dump_2d_plt_file_single(filename, np_density, np_u, np_v, B, C):
    """Tecplot POINT zone: X Y RHO U V from staggered u, v (utils/io.py:336-352)."""
    shape = np_density.shape
    np_u = 0.5 * (np_u[..., 1:] + np_u[..., :-1])
    np_v = 0.5 * (np_v[..., 1:, :] + np_v[..., :-1, :])
    _write_zone(filename, '"X" "Y" "RHO" "U" "V" ', shape,
                ((i, j, np_density[B, C, j, i], np_u[B, C, j, i], np_v[B, C, j, i])
                 for j in range(shape[-2]) for i in range(shape[-1])))
