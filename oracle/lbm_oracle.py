"""CPU oracle for the D2Q9 lattice-update path of GeCao/JRL-LBM.  TEST INFRASTRUCTURE ONLY.

A numpy restatement of what the reference computes per step, function by function, each
citing the reference file:line it follows (paths relative to /root/reference/).  It is the
checker for the CUDA path in jrl-lbm_b200/; only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import it.  The product never does.

Parity status: PINNED.  The reference ships no tests or golden vectors,
so the pins are outputs of the reference itself run in the build container:
  * every kernel restated here is checked against the reference's own kernel body compiled
    for the CPU (oracle/_ref/libjrl_ref.so, built by oracle/build_ref.sh from the .cu files
    where they lie) -- tests/test_oracle_vs_ref.py, skipped where that .so is absent;
  * every torch-op path restated here (feq, BGK/MRT/KBC collision, moments, contact angle,
    axisymmetric source) is checked against tests/golden/*.npz, written by
    oracle/gen_golden.py from the reference's unmodified Python classes imported on the CPU
    (oracle/ref_cpu.py) -- tests/test_oracle_golden.py.
  * the torch-op paths follow torch's CUDA arithmetic, because the reference hard-requires CUDA
    (readme.md:7): division by a python scalar = multiplication by the fp32 reciprocal, `sum(dim=1)`
    = four interleaved accumulators, batched 9x9 mat-vec = two FMA chains (`_sum9`, `_matvec9`).
    Those orders were measured on a B200 against the reference-as-run (tools/bench_ref_cuda.py,
    tools/exp_torch_order.py) and are pinned by tests/golden/refcuda_*.npz = outputs of the
    reference's own fast-math CUDA build -- tests/test_oracle_golden.py::test_oracle_vs_refcuda.
    (The CPU-torch fixtures above differ from them in the last bit; the 1e-5 bar covers both.)
One documented deviation: OUTFLOW cells (propagation.cu:195-255) read a neighbour's f_new
that another CUDA thread writes in the same launch (a data race).  The oracle, like the new
kernels, defines the result as "the neighbour's value after streaming and inflow reset".

Conventions: f/h [B,9,H,W], rho/C/phi [B,1,H,W], vel/force/mesh [B,2,H,W]
(channel 0 = x), flags [B,1,H,W] uint8, all C-contiguous float32.  Arithmetic type is float32
with float64 intermediates exactly where the reference's C++ literals force them.
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32
F64 = np.float64

# --- R0: lattice and flag constants (propagation.cu:9-20, 49-64; utils/types.py:5-15) ----------
NOTHING, FLUID, OBSTACLE, EMPTY, INFLOW, OUTFLOW, INFLOW_2 = 0, 1, 2, 4, 8, 16, 32
AXI_NOT, AXI_X, AXI_Y = 0, 1, 2
KBC_LBGK, KBC_A, KBC_B, KBC_C, KBC_D = 0, 0b10000101, 0b10000110, 0b10001001, 0b10001010

E = np.array([[0, 0], [1, 0], [0, 1], [-1, 0], [0, -1], [1, 1], [-1, 1], [-1, -1], [1, -1]], dtype=np.int64)
WEIGHT = np.array([4 / 9] + [1 / 9] * 4 + [1 / 36] * 4, dtype=F32)
OPP = np.array([0, 3, 4, 1, 2, 7, 8, 5, 6])
MIRROR_X = np.array([0, 3, 2, 1, 4, 6, 5, 8, 7])  # e_idx_x_axisymmetric
MIRROR_Y = np.array([0, 1, 4, 3, 2, 8, 7, 6, 5])  # e_idx_y_axisymmetric


def _sdiv(a, s):
    """tensor / python-scalar the way torch evaluates it: a * float32(1/s)
    (ATen div_true_kernel multiplies by the reciprocal when the divisor is a CPU scalar)."""
    return a * F32(1.0 / s)


def meshgrid(B, H, W, dtype=F32):
    """utils/grid.py:67-82 -- [B,2,H,W], channel 0 = x index, channel 1 = y index."""
    y, x = np.meshgrid(np.arange(H, dtype=dtype), np.arange(W, dtype=dtype), indexing="ij")
    return np.broadcast_to(np.stack([x, y])[None], (B, 2, H, W)).copy()


def union_phi_obs(p1, p2, alpha=0.0):
    """utils/functions.py:4-15."""
    if alpha == 1.0:
        return np.minimum(p1, p2)
    return F32(1.0 / (1.0 + alpha)) * (p1 + p2 - np.sqrt(p1 * p1 + p2 * p2 - 2 * alpha * p1 * p2))


# --- neighbour index rules ---------------------------------------------------------------------

def _neighbours(flags, axisym, mode):
    """Per-cell x_neg/x_pos/y_neg/y_pos [B,H,W] int arrays.
    mode 'mirror': streaming rule, propagation.cu:74-87 (FLUID wraps, others mirror 1 / W-2).
    mode 'clamp' : gradient / wall rule, collision_2d.cu:59-72, solid_fluid_coupling.cu:98-112."""
    B, _, H, W = flags.shape
    fl = flags[:, 0]
    y, x = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    y = np.broadcast_to(y, (B, H, W))
    x = np.broadcast_to(x, (B, H, W))
    fluid = fl == FLUID
    on_x = (x == 0) & (axisym == AXI_X)
    on_y = (y == 0) & (axisym == AXI_Y)
    if mode == "mirror":
        xn_o = np.where(x - 1 < 0, 1, x - 1)
        xp_o = np.where(x + 1 >= W, W - 2, x + 1)
        yn_o = np.where(y - 1 < 0, 1, y - 1)
        yp_o = np.where(y + 1 >= H, H - 2, y + 1)
    else:
        xn_o = np.maximum(0, x - 1)
        xp_o = np.minimum(W - 1, x + 1)
        yn_o = np.maximum(0, y - 1)
        yp_o = np.minimum(H - 1, y + 1)
    xn = np.where(on_x, 0, np.where(fluid, (x - 1 + W) % W, xn_o))
    xp = np.where(fluid, (x + 1) % W, xp_o)
    yn = np.where(on_y, 0, np.where(fluid, (y - 1 + H) % H, yn_o))
    yp = np.where(fluid, (y + 1) % H, yp_o)
    return y, x, xn, xp, yn, yp, on_x, on_y


def _income_outcome(y, x, xn, xp, yn, yp):
    """Upstream ('income') and downstream ('outcome') cell coordinates per direction,
    propagation.cu:88-108."""
    inc = [(y, x), (y, xn), (yn, x), (y, xp), (yp, x), (yn, xn), (yn, xp), (yp, xp), (yp, xn)]
    out = [(y, x), (y, xp), (yp, x), (y, xn), (yn, x), (yp, xp), (yp, xn), (yn, xn), (yn, xp)]
    return inc, out


def _equilibrium_mixed(rho_t, ux, uy):
    """Inflow equilibrium with the reference kernel's mixed precision, propagation.cu:340-350:
    cs2, uv, eu are float; `1.0 + eu / cs2` and the whole second term are double; each of the
    two statements rounds to float on store."""
    cs2 = F32(1.0 / 3.0)
    uv = ux * ux + uy * uy
    out = []
    for q in range(9):
        eu = ux * F32(E[q, 0]) + uy * F32(E[q, 1])
        rw = rho_t * WEIGHT[q]
        first = (rw.astype(F64) * (1.0 + (eu / cs2).astype(F64))).astype(F32)
        second = rw.astype(F64) * (0.5 * eu.astype(F64) * eu.astype(F64) / F64(cs2) / F64(cs2)
                                   - 0.5 * uv.astype(F64) / F64(cs2))
        out.append((first.astype(F64) + second).astype(F32))
    return out


# --- R1: streaming with all link rules (K1) ----------------------------------------------------

def stream(flags, f, phi_obs=None, rho=None, vel=None, inflow_vel=None, inflow_density=-1.0, axisym=0):
    """kernel_prop_2d_forward, propagation.cu:22-377, with the wrapper's conventions
    (LBM_propagation_2d.py:35-92).  rho and vel are MUTATED in place at INFLOW / INFLOW_2 cells
    (propagation.cu:300, 320-321, 362).  Returns f_new."""
    B, Q, H, W = f.shape
    assert Q == 9 and flags.dtype == np.uint8 and f.dtype == F32
    fl = flags[:, 0]
    y, x, xn, xp, yn, yp, on_x, on_y = _neighbours(flags, axisym, "mirror")
    inc, out = _income_outcome(y, x, xn, xp, yn, yp)
    b = np.broadcast_to(np.arange(B)[:, None, None], (B, H, W))
    self_obs = fl == OBSTACLE
    f_new = np.empty_like(f)
    f_new[:, 0] = f[:, 0]                                                # :110
    t = None if phi_obs is None else phi_obs[:, 0]
    for q in range(1, 9):
        qi = np.full((B, H, W), q)                                        # income_Q_idx, :116-124 / :135-143
        if q in (5, 8, 1):
            qi = np.where(on_x, MIRROR_X[q], qi)
        if q in (5, 6, 2):
            qi = np.where(on_y & ~on_x, MIRROR_Y[q], qi)
        iy, ix = inc[q]
        pulled = f[b, qi, iy, ix]
        # self OBSTACLE: store the reflected incoming set, :111-128
        f_new[:, OPP[q]][self_obs] = pulled[self_obs]
        up_obs = fl[b, iy, ix] == OBSTACLE                                # :145
        f_this = f[:, OPP[q]]                                             # :154
        if t is None:
            bounced = f_this                                              # :155-159
        else:
            t1 = 2 * t - 1                                                # :166
            near = t1 * pulled + (1 - t1) * f_this                        # :168
            qo = np.full((B, H, W), q)                                    # outcome_Q_idx, :172-180
            if q in (6, 7, 3):
                qo = np.where(on_x, MIRROR_X[q], qo)
            if q in (7, 8, 4):
                qo = np.where(on_y & ~on_x, MIRROR_Y[q], qo)
            oy, ox = out[q]
            f_next = f[b, OPP[qo], oy, ox]                                # :182
            t2 = 2 * t
            far = t2 * f_this + (1 - t2) * f_next                         # :183
            with np.errstate(invalid="ignore"):
                bounced = np.where((t >= 0.5) & (t <= 1), near, np.where((t <= 0.5) & (t >= 0), far, f_this))
        val = np.where(up_obs, bounced, pulled)
        f_new[:, q][~self_obs] = val[~self_obs]

    # INFLOW, :295-352
    m = fl == INFLOW
    if m.any():
        if rho is not None and inflow_density >= 0:
            rho[:, 0][m] = F32(inflow_density)
            rho_t = np.full(int(m.sum()), inflow_density, F32)
        elif rho is not None:
            rho_t = rho[:, 0][m]
        elif inflow_density >= 0:
            rho_t = np.full(int(m.sum()), inflow_density, F32)
        else:
            rho_t = np.zeros(int(m.sum()), F32)
        if vel is not None and inflow_vel is not None:
            vel[:, 0][m] = inflow_vel[0]
            vel[:, 1][m] = inflow_vel[1]
            ux = np.full(int(m.sum()), inflow_vel[0], F32)
            uy = np.full(int(m.sum()), inflow_vel[1], F32)
        elif vel is not None:
            ux, uy = vel[:, 0][m], vel[:, 1][m]
        elif inflow_vel is not None:
            ux = np.full(int(m.sum()), inflow_vel[0], F32)
            uy = np.full(int(m.sum()), inflow_vel[1], F32)
        else:
            ux = uy = np.zeros(int(m.sum()), F32)
        for q, v in enumerate(_equilibrium_mixed(rho_t, ux, uy)):
            f_new[:, q][m] = v
    # INFLOW_2, :354-377
    m = fl == INFLOW_2
    if m.any():
        n = int(m.sum())
        ux = np.zeros(n, F32) if vel is None else vel[:, 0][m]
        uy = np.zeros(n, F32) if vel is None else vel[:, 1][m]
        if rho is not None:
            rho[:, 0][m] = F32(1)
        for q, v in enumerate(_equilibrium_mixed(np.ones(n, F32), ux, uy)):
            f_new[:, q][m] = v
    # OUTFLOW, :195-255 (deterministic reading of the racy copy, see module docstring)
    m = fl == OUTFLOW
    if m.any():
        sy = np.where(y == 0, y + 1, np.where(y == H - 1, y - 1, y))
        sx = np.where(x == 0, x + 1, np.where(x == W - 1, x - 1, x))
        src = f_new[b[m], :, sy[m], sx[m]].copy()                         # [n, 9]
        for q in range(9):
            f_new[:, q][m] = src[:, q]
    return f_new


def rebounce_obstacle(f, flags):
    """LBM_propagation_2d.py:181-186."""
    return np.where(flags == OBSTACLE, f[:, OPP], f)


# --- R6 stencils: K2 / K3 ----------------------------------------------------------------------

def _clamped_outcome(flags, axisym):
    y, x, xn, xp, yn, yp, on_x, on_y = _neighbours(flags, axisym, "clamp")
    _, out = _income_outcome(y, x, xn, xp, yn, yp)
    inc, _ = _income_outcome(y, x, xn, xp, yn, yp)
    return y, x, inc, out, on_x, on_y


def grad(s, flags, dx=1.0, axisym=0):
    """kernel_get_grad_2d_forward, collision_2d.cu:22-148.  s [B,1,H,W] -> [B,2,H,W]."""
    B, _, H, W = s.shape
    fl = flags[:, 0]
    _, _, _, out, _, _ = _clamped_outcome(flags, axisym)
    b = np.broadcast_to(np.arange(B)[:, None, None], (B, H, W))
    v = [s[:, 0][b, oy, ox] for (oy, ox) in out]
    ob = [fl[b, oy, ox] == OBSTACLE for (oy, ox) in out]
    gx = np.zeros((B, H, W), F32)
    gy = np.zeros((B, H, W), F32)
    gx = gx + np.where(~ob[1] & ~ob[3], 4 * (v[1] - v[3]), F32(0))
    gx = gx + np.where(~ob[5] & ~ob[6], v[5] - v[6], F32(0))
    gx = gx + np.where(~ob[7] & ~ob[8], v[8] - v[7], F32(0))
    gx = np.where(ob[1] | ob[3] | ob[5] | ob[6] | ob[7] | ob[8], F32(0), gx)
    gy = gy + np.where(~ob[2] & ~ob[4], 4 * (v[2] - v[4]), F32(0))
    gy = gy + np.where(~ob[5] & ~ob[8], v[5] - v[8], F32(0))
    gy = gy + np.where(~ob[6] & ~ob[7], v[6] - v[7], F32(0))
    gy = np.where(ob[2] | ob[4] | ob[5] | ob[8] | ob[6] | ob[7], F32(0), gy)
    g = np.stack([gx, gy], axis=1)
    g = (g.astype(F64) / (12 * float(dx))).astype(F32)                    # `/= 12 * dx`, dx is double
    g[np.broadcast_to((fl == OBSTACLE)[:, None], g.shape)] = 0
    return g


def div(v2, flags, dx=1.0, axisym=0):
    """kernel_get_div_2d_forward, collision_2d.cu:150-220.  v2 [B,2,H,W] -> [B,1,H,W]."""
    B, _, H, W = v2.shape
    fl = flags[:, 0]
    _, _, _, out, _, _ = _clamped_outcome(flags, axisym)
    b = np.broadcast_to(np.arange(B)[:, None, None], (B, H, W))
    a = [v2[:, 0][b, oy, ox] for (oy, ox) in out]
    c = [v2[:, 1][b, oy, ox] for (oy, ox) in out]
    d = np.zeros((B, H, W), F32)
    d = d + 4 * (a[1] - a[3])
    d = d + (a[5] - a[6])
    d = d + (a[8] - a[7])
    d = d + 4 * (c[2] - c[4])
    d = d + (c[5] - c[8])
    d = d + (c[6] - c[7])
    d = (d.astype(F64) / (12 * float(dx))).astype(F32)
    d[fl == OBSTACLE] = 0
    return d[:, None]


# --- R5: regularised moving-wall reconstruction (K7) --------------------------------------------

def wall(dt, dx, tau, vel, flags, f, phi_obs, obs_vel=None, axisym=0):
    """kernel_2d_forward, solid_fluid_coupling.cu:25-252 (arguments `rho`, `is_convection`
    are unused there).  Returns a new f."""
    B, _, H, W = f.shape
    fl = flags[:, 0]
    y, x, inc, out, on_x, on_y = _clamped_outcome(flags, axisym)
    f_new = f.copy()
    cand = (fl != OBSTACLE) & (x != W - 1) & (y != H - 1)                 # :85-96
    bb, yy, xx = np.nonzero(cand)
    if bb.size == 0:
        return f_new
    c = F32(dx / dt)
    cs2 = F32(F64(c * c) / 3.0)
    eps = F32(1e-10)
    t = phi_obs[:, 0][bb, yy, xx]
    n = bb.size
    vt = [np.zeros(n, F32), np.zeros(n, F32)]
    rho_t = np.zeros(n, F32)
    is_obs = np.zeros((9, n), bool)
    count = np.zeros(n, np.int64)
    with np.errstate(invalid="ignore"):
        t_ok = (t > -eps) & (t <= 1 + eps)                                # :152
    for i in range(9):
        inv = OPP[i]
        oy, ox = out[i][0][bb, yy, xx], out[i][1][bb, yy, xx]
        hit = (fl[bb, oy, ox] == OBSTACLE) & t_ok
        is_obs[i] = hit
        count += hit
        iy, ix = inc[i][0][bb, yy, xx], inc[i][1][bb, yy, xx]
        ux = vel[:, 0][bb, iy, ix].copy()
        uy = vel[:, 1][bb, iy, ix].copy()
        if i in (1, 5, 8):
            ux = np.where(on_x[bb, yy, xx], -ux, ux)                       # :160-163
        if i in (2, 5, 6):
            uy = np.where(on_y[bb, yy, xx], -uy, uy)
        tp1 = t.astype(F64) + 1.0
        if obs_vel is not None:
            ox_u = obs_vel[:, 0][bb, oy, ox]
            oy_u = obs_vel[:, 1][bb, oy, ox]
            ax = (vt[0].astype(F64) + (t * ux + ox_u).astype(F64) / tp1).astype(F32)   # :172-173
            ay = (vt[1].astype(F64) + (t * uy + oy_u).astype(F64) / tp1).astype(F32)
            dot = F32(E[inv, 0]) * ox_u + F32(E[inv, 1]) * oy_u
            ar = (rho_t.astype(F64) + 6.0 * F64(WEIGHT[i]) * dot.astype(F64)).astype(F32)  # :176
        else:
            ax = (vt[0].astype(F64) + (t * ux).astype(F64) / tp1).astype(F32)          # :180-181
            ay = (vt[1].astype(F64) + (t * uy).astype(F64) / tp1).astype(F32)
            ar = rho_t
        vt[0] = np.where(hit, ax, vt[0])
        vt[1] = np.where(hit, ay, vt[1])
        rho_t = np.where(hit, ar, rho_t)
        rho_t = rho_t + f[bb, inv, yy, xx]                                 # :198
    act = count > 0
    if not act.any():
        return f_new
    cnt = np.where(act, count, 1).astype(F32)
    vt = [vt[0] / cnt, vt[1] / cnt]                                        # :205-206

    def g(comp, pa, pb, d1a, d1b, d2a, d2b):
        u = vel[:, comp]
        s = lambda k: u[bb, out[k][0][bb, yy, xx], out[k][1][bb, yy, xx]].astype(F64)
        return ((4.0 * s(pa) - 4.0 * s(pb) + s(d1a) - s(d1b) + s(d2a) - s(d2b)) / 12.0).astype(F32)

    g00 = g(0, 1, 3, 5, 6, 8, 7)                                           # :210-237
    g01 = g(0, 2, 4, 5, 8, 6, 7)
    g10 = g(1, 1, 3, 5, 6, 8, 7)
    g11 = g(1, 2, 4, 5, 8, 6, 7)
    tau = float(tau)
    rc = (rho_t * cs2).astype(F64) * tau
    for i in range(9):
        inv = OPP[i]
        ex, ey = int(E[inv, 0]), int(E[inv, 1])
        pab = np.zeros(n, F32)
        terms = [
            ((rho_t * vt[0] * vt[0]).astype(F64) - rc * (g00 + g00).astype(F64), F32(ex * ex) - cs2),
            ((rho_t * vt[1] * vt[0]).astype(F64) - rc * (g10 + g01).astype(F64), F32(ey * ex) - F32(0)),
            ((rho_t * vt[0] * vt[1]).astype(F64) - rc * (g01 + g10).astype(F64), F32(ex * ey) - F32(0)),
            ((rho_t * vt[1] * vt[1]).astype(F64) - rc * (g11 + g11).astype(F64), F32(ey * ey) - cs2),
        ]
        for a, bq in terms:                                                # :244-247
            pab = (pab.astype(F64) + a * F64(bq)).astype(F32)
        lin = rho_t + rho_t * c / cs2 * (vt[0] * F32(ex) + vt[1] * F32(ey))
        val = (F64(WEIGHT[inv]) * (lin.astype(F64) + 0.5 * pab.astype(F64) / F64(cs2) / F64(cs2))).astype(F32)  # :248-249
        sel = is_obs[i] & act
        f_new[bb[sel], inv, yy[sel], xx[sel]] = val[sel]
    return f_new


# --- R2 / R7: moments ---------------------------------------------------------------------------

def contact_angle_correctness(rho, flags, contact_angle):
    """LBM_macro_compute_2d.py:237-271."""
    ro = np.zeros_like(rho)
    tn = F32(math.tan(math.pi / 2.0 - float(contact_angle)))
    if isinstance(contact_angle, (np.ndarray, np.generic)):
        tn = np.tan(F32(math.pi / 2.0) - F32(contact_angle)).astype(F32)  # tensor contact angle: fp32 tan
    ro[..., 1:-1, 0] = rho[..., 1:-1, 2] + tn * np.abs(rho[..., 2:, 1] - rho[..., :-2, 1])
    ro[..., 1:-1, -1] = rho[..., 1:-1, -3] + tn * np.abs(rho[..., 2:, -2] - rho[..., :-2, -2])
    ro[..., 0, 1:-1] = rho[..., 2, 1:-1] + tn * np.abs(rho[..., 1, 2:] - rho[..., 1, :-2])
    ro[..., -1, 1:-1] = rho[..., -3, 1:-1] + tn * np.abs(rho[..., -2, 2:] - rho[..., -2, :-2])
    ro[..., 0, 0] = 0.5 * (ro[..., 1, 0] + ro[..., 0, 1])
    ro[..., -1, 0] = 0.5 * (ro[..., -2, 0] + ro[..., -1, 1])
    ro[..., 0, -1] = 0.5 * (ro[..., 0, -2] + ro[..., 1, -1])
    ro[..., -1, -1] = 0.5 * (ro[..., -2, -2] + ro[..., -2, -2])           # sic (:269)
    return np.where(flags == OBSTACLE, ro, rho)


def _sum9(f):
    """`t.sum(dim=1, keepdim=True)` as torch evaluates it on CUDA for a 9-long reduction: one thread per output,
    vt0 = 4 interleaved accumulators (element k into accumulator k % 4), combined in order
    (ATen/native/cuda/Reduce.cuh thread_reduce_impl): ((((a0+a4)+a8) + (a1+a5)) + (a2+a6)) + (a3+a7)."""
    a = [f[:, q] for q in range(9)]
    s = (((a[0] + a[4]) + a[8]) + (a[1] + a[5])) + (a[2] + a[6])
    return (s + (a[3] + a[7]))[:, None]


def _fma(a, b, c):
    """fmaf(a, b, c) for float32 arrays: the product is exact in float64; the sum is rounded to float64 and then
    to float32 (double rounding differs from a true FMA with probability ~2^-29 per operation)."""
    return (np.asarray(a, F64) * np.asarray(b, F64) + np.asarray(c, F64)).astype(F32)


def _dot9(coef, x):
    """One row of torch-CUDA's batched 9x9 mat-vec: k = 0..4 and k = 5..8 as two FMA chains from 0, halves added
    (cuBLAS kernel for this shape; measured by tools/exp_torch_order.py).  coef: 9 scalars or per-cell arrays."""
    lo = np.zeros_like(x[:, 0])
    hi = np.zeros_like(x[:, 0])
    for k in range(5):
        lo = _fma(coef[k], x[:, k], lo)
    for k in range(5, 9):
        hi = _fma(coef[k], x[:, k], hi)
    return lo + hi


def macro_compute(dx, dt, f, rho, vel, flags, contact_angle=None):
    """LBMMacroCompute2d.macro_compute, LBM_macro_compute_2d.py:316-351 (density=None branch)."""
    c = dx / dt
    obs = flags == OBSTACLE
    rho_new = np.where(obs, rho, _sum9(f))
    if contact_angle is not None:
        rho_new = contact_angle_correctness(rho_new, flags, contact_angle)
    mx = _sum9(f * E[:, 0].astype(F32).reshape(1, 9, 1, 1))[:, 0]      # (f.unsqueeze(2) * e).sum(1), :335
    my = _sum9(f * E[:, 1].astype(F32).reshape(1, 9, 1, 1))[:, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = F32(c) / rho_new
        vel_new = np.stack([mx, my], axis=1) * inv
    vel_new = np.where(obs, vel, vel_new)
    return rho_new.astype(F32), vel_new.astype(F32)


def macro_compute_kernel(dx, dt, f, rho, vel, flags):
    """kernel_macompute_2d_forward (ApplyCompute_fluid), macro_compute.cu:80-157."""
    c = F32(dx / dt)
    obs = flags == OBSTACLE
    r = np.zeros_like(f[:, 0])
    for q in range(9):
        r = r + f[:, q]
    mx = np.zeros_like(r)
    my = np.zeros_like(r)
    for q in range(9):
        mx = mx + f[:, q] * F32(E[q, 0]) * c
        my = my + f[:, q] * F32(E[q, 1]) * c
    with np.errstate(divide="ignore", invalid="ignore"):
        v = np.stack([mx / r, my / r], axis=1)
    return np.where(obs, rho, r[:, None]), np.where(obs, vel, v)


def macro_compute_C(h, C, flags):
    """LBMMacroCompute2d.macro_compute_C, LBM_macro_compute_2d.py:273-295."""
    return np.where(flags == OBSTACLE, C, _sum9(h))


def macro_compute_C_kernel(dx, dt, h, C, flags, vel, mesh_grid, axisym):
    """kernel_macompute_C_2d_forward (ApplyCompute_C), macro_compute.cu:22-78."""
    s = np.zeros_like(h[:, 0])
    for q in range(9):
        s = s + h[:, q]
    if axisym == AXI_X:
        ur, r = vel[:, 0], mesh_grid[:, 0]
        with np.errstate(divide="ignore", invalid="ignore"):
            den = 1.0 + (0.5 * float(dt) * ur.astype(F64)) / r.astype(F64)   # :75, double
            s = (s.astype(F64) / den).astype(F32)
    return np.where(flags == OBSTACLE, C, s[:, None])


def get_vort(vel, dx):
    """LBM_macro_compute_2d.py:353-361."""
    v = _sdiv((vel[:, 0:1, 2:, 1:-1] - vel[:, 0:1, :-2, 1:-1]) - (vel[:, 1:2, 1:-1, 2:] - vel[:, 1:2, 1:-1, :-2]), 2.0 * dx)
    return np.pad(v, ((0, 0), (0, 0), (1, 1), (1, 1)), mode="edge")


# --- R3: equilibrium ----------------------------------------------------------------------------

def get_feq(dx, dt, rho, vel, force=None):
    """LBMCollision2d.get_feq_static_, LBM_collision_2d.py:236-263."""
    if force is not None:
        with np.errstate(divide="ignore", invalid="ignore"):
            vel = vel + np.where(rho <= 1e-10, F32(0), 0.5 * force / rho)
    c = dx / dt
    cs2 = c * c / 3.0
    uv = (vel[:, 0] * vel[:, 0] + vel[:, 1] * vel[:, 1])[:, None]
    feq = np.empty((rho.shape[0], 9) + rho.shape[2:], F32)
    for q in range(9):
        eu = (vel[:, 0] * F32(E[q, 0]) * F32(c) + vel[:, 1] * F32(E[q, 1]) * F32(c))[:, None]
        rw = rho * WEIGHT[q]
        a = rw * (1.0 + _sdiv(eu, cs2))
        bq = rw * (_sdiv(_sdiv(0.5 * eu * eu, cs2), cs2) - _sdiv(0.5 * uv, cs2))
        feq[:, q:q + 1] = a + bq
    return feq


# --- R6: axisymmetric source --------------------------------------------------------------------

def axisym_source(dx, dt, rho, vel, flags, feq, tau, axisym, is_convection, mesh_grid):
    """get_G_source_from_axisymmetric, LBM_collision_2d.py:343-471. `tau` is tau_D when
    is_convection.  Returns [B,9,H,W] or the python float 0.0."""
    c = dx / dt
    cs2 = c * c / 3.0
    ef = E.astype(F32)
    if axisym == AXI_X:
        r = mesh_grid[:, 0:1]
        er = ef[:, 0].reshape(1, 9, 1, 1)
        ur = vel[:, 0:1]
        A1 = -rho * ur / r
        if is_convection:
            s = (1 - 0.5 / tau) * er / r
            return -ur / r * feq * (1 - 0.5 / tau - 0.5 * s)              # :383-384 (no weight factor)
        visc = cs2 * (tau - 0.5)
        mu = visc * rho
        pressure = rho * F32(cs2)
        gu = grad(np.ascontiguousarray(vel[:, 0:1]), flags, dx, axisym)
        gv = grad(np.ascontiguousarray(vel[:, 1:2]), flags, dx, axisym)
        A2 = grad(pressure, flags, dx, axisym)[:, 0:1]
        A2 = A2 + div(rho * ur * vel, flags, dx, axisym)
        A2 = A2 * (dt / (2 * r))
        F2 = (mu / r) * np.concatenate((gu[:, 0:1] - ur / r, gv[:, 0:1]), axis=1)
        F2 = F2 + vel * A1
        F2 = F2 - dt * (tau - 1) * cs2 * grad(A1.astype(F32), flags, dx, axisym)
        Fe = np.zeros((rho.shape[0], 9) + rho.shape[2:], F32)
        for q in range(9):
            Fe[:, q] = F2[:, 0] * ef[q, 0] + F2[:, 1] * ef[q, 1]
        G = A1 + A2 + _sdiv(Fe, cs2)
        return (G * WEIGHT.reshape(1, 9, 1, 1)).astype(F32)
    if axisym == AXI_Y:
        r = mesh_grid[:, 1:2] - F32(0.5 * dx)
        r = r.copy()
        r[..., 0, :] = 1
        ur = vel[:, 1:2]
        A1 = -rho * ur / r
        if is_convection:
            G = A1 * np.ones((1, 1, 1, 1), F32)
            return (G * WEIGHT.reshape(1, 9, 1, 1)).astype(F32)
        visc = cs2 * (tau - 0.5)
        mu = visc * rho
        gu = grad(np.ascontiguousarray(vel[:, 0:1]), flags, dx, axisym)
        gv = grad(np.ascontiguousarray(vel[:, 1:2]), flags, dx, axisym)
        H1 = mu * (gu[:, 1:2] + gv[:, 0:1]) / r - rho * vel[:, 0:1] * vel[:, 1:2] / r
        H2 = mu * (gv[:, 1:2] * 2) / r - 2 * mu * vel[:, 1:2] / (r * r) - rho * vel[:, 1:2] * vel[:, 1:2] / r
        ex = ef[:, 0].reshape(1, 9, 1, 1)
        ey = ef[:, 1].reshape(1, 9, 1, 1)
        G = A1 + _sdiv(ex * H1 + ey * H2, cs2)
        return (G * WEIGHT.reshape(1, 9, 1, 1)).astype(F32)
    return 0.0


# --- R4 / R4m / R4k: collision operators --------------------------------------------------------

def collide_bgk(dx, dt, f, rho, vel, flags, tau, force=None, mesh_grid=None, is_convection=False,
                tau_D=None, axisym=0):
    """LBMCollision2d.collision, LBM_collision_2d.py:473-555."""
    t = tau_D if is_convection else tau
    feq = get_feq(dx, dt, rho, vel, force)
    if is_convection:
        s = 0.0
        if axisym == AXI_X:                                               # :516-521 (the y-axis twin at :522 is dead)
            er = E[:, 0].astype(F32).reshape(1, 9, 1, 1)
            r = mesh_grid[:, 0:1]
            s = dt * (1 - 0.5 / t) * er / r
        w = 1.0 / t + s
    else:
        w = 1.0 / t
    Gi = axisym_source(dx, dt, rho, vel, flags, feq, t, axisym, is_convection, mesh_grid)
    out = f + w * (feq - f) + dt * Gi
    return np.where(flags == OBSTACLE, f, out).astype(F32)


MRT_M = np.array([
    [1, 1, 1, 1, 1, 1, 1, 1, 1],
    [-4, -1, -1, -1, -1, 2, 2, 2, 2],
    [4, -2, -2, -2, -2, 1, 1, 1, 1],
    [0, 1, 0, -1, 0, 1, -1, -1, 1],
    [0, -2, 0, 2, 0, 1, -1, -1, 1],
    [0, 0, 1, 0, -1, 1, 1, -1, -1],
    [0, 0, -2, 0, 2, 1, 1, -1, -1],
    [0, 1, -1, 1, -1, 0, 0, 0, 0],
    [0, 0, 0, 0, 0, 1, -1, 1, -1],
], dtype=F32)                                                             # LBM_collision_MRT_2d.py:41-57


def mrt_matrix(tau):
    """M^-1 S M, LBM_collision_MRT_2d.py:59-69 (fp32 like the reference)."""
    s7 = 1.0 / tau
    s4 = 8.0 * (2 - s7) / (8 - s7)
    S = np.array([0.0, s7, s7, 0.0, s4, 0.0, s4, s7, s7], dtype=F32).reshape(9, 1)
    Minv = np.linalg.inv(MRT_M).astype(F32)
    return (Minv @ (S * MRT_M)).astype(F32)


def collide_mrt(dx, dt, f, rho, vel, flags, tau, force=None, mesh_grid=None, axisym=0, mat=None):
    """LBMCollisionMRT2d.collision, LBM_collision_MRT_2d.py:84-102, 121-184."""
    feq = get_feq(dx, dt, rho, vel, force)
    mat = mrt_matrix(tau) if mat is None else mat
    d = feq - f
    omega = np.zeros_like(f)
    for i in range(9):
        omega[:, i] = _dot9(mat[i], d)
    Gi = axisym_source(dx, dt, rho, vel, flags, feq, tau, axisym, False, mesh_grid)
    out = f + omega + dt * Gi
    return np.where(flags == OBSTACLE, f, out).astype(F32)


KBC_CMAT = np.array([
    [1, 1, 1, 1, 1, 1, 1, 1, 1],
    [0, 1, 0, -1, 0, 1, -1, -1, 1],
    [0, 0, 1, 0, -1, 1, 1, -1, -1],
    [0, 1, 0, 1, 0, 1, 1, 1, 1],
    [0, 0, 0, 0, 0, 1, -1, 1, -1],
    [0, 0, 1, 0, 1, 1, 1, 1, 1],
    [0, 0, 0, 0, 0, 1, 1, -1, -1],
    [0, 0, 0, 0, 0, 1, -1, -1, 1],
    [0, 0, 0, 0, 0, 1, 1, 1, 1],
], dtype=F32)                                                             # LBM_collision_KBC_2d.py:38-55
KBC_P = [0, 1, 0, 2, 1, 0, 2, 1, 2]
KBC_Q = [0, 0, 1, 0, 1, 2, 1, 2, 2]


def kbc_omega(rho, vel, f, feq, tau, kbc_type):
    """LBMCollisionKBC2d.get_Omega, LBM_collision_KBC_2d.py:70-249."""
    if kbc_type in (KBC_A, KBC_B):
        dxm = [F32(E[q, 0]) - vel[:, 0] for q in range(9)]               # (c e - u) with c = 1, :96-105
        dym = [F32(E[q, 1]) - vel[:, 1] for q in range(9)]

        def mom(g, k):
            return _dot9([(dxm[q] ** KBC_P[k]) * (dym[q] ** KBC_Q[k]) for q in range(9)], g)[:, None]
    else:
        def mom(g, k):
            return _dot9(KBC_CMAT[k], g)[:, None]
    M20, M11, M02 = mom(f, 3), mom(f, 4), mom(f, 5)
    Q20, Q11, Q02 = mom(feq, 3), mom(feq, 4), mom(feq, 5)
    T, N, P = M20 + M02, M20 - M02, M11
    Te, Ne, Pe = Q20 + Q02, Q20 - Q02, Q11
    df = f - feq
    ux, uy = vel[:, 0:1], vel[:, 1:2]
    if kbc_type == KBC_LBGK:
        ds, dh = df, np.zeros_like(f)
    elif kbc_type in (KBC_A, KBC_C):
        def shear(T, N, P):
            return [rho - T, 0.5 * (0.5 * (T + N) + rho * ux), 0.5 * (0.5 * (T - N) + rho * uy),
                    0.5 * (0.5 * (T + N) - rho * ux), 0.5 * (0.5 * (T - N) - rho * uy),
                    0.25 * P, 0.25 * (-P), 0.25 * P, 0.25 * (-P)]
        ds = np.concatenate([a - b for a, b in zip(shear(T, N, P), shear(Te, Ne, Pe))], axis=1)
        dh = df - ds
    elif kbc_type in (KBC_B, KBC_D):
        def shear(N, P):
            return [rho + 0 * N, 0.5 * (0.5 * N + rho * ux), 0.5 * (0.5 * (-N) + rho * uy),
                    0.5 * (0.5 * N - rho * ux), 0.5 * (0.5 * (-N) - rho * uy),
                    0.25 * P, 0.25 * (-P), 0.25 * P, 0.25 * (-P)]
        ds = np.concatenate([a - b for a, b in zip(shear(N, P), shear(Ne, Pe))], axis=1)
        dh = df - ds
    else:
        raise RuntimeError("Please specific a right KBC Type value")
    beta = 1.0 / (2.0 * tau)
    gamma = 2.0
    eps = 1e-10
    if kbc_type != KBC_LBGK:
        with np.errstate(divide="ignore", invalid="ignore"):
            dsdh = _sum9(ds * dh / (feq + eps))
            dhdh = _sum9(dh * dh / (feq + eps))
            gamma = 1.0 / beta - (2.0 - 1.0 / beta) * (dsdh / (dhdh + eps))
    return (-beta * (2.0 * ds + gamma * dh)).astype(F32)


def collide_kbc(dx, dt, f, rho, vel, flags, tau, kbc_type, force=None, mesh_grid=None, axisym=0):
    """LBMCollisionKBC2d.collision, LBM_collision_KBC_2d.py:251-317."""
    feq = get_feq(dx, dt, rho, vel, force)
    Gi = axisym_source(dx, dt, rho, vel, flags, feq, tau, axisym, False, mesh_grid)
    out = f + kbc_omega(rho, vel, f, feq, tau, kbc_type) + dt * Gi
    return np.where(flags == OBSTACLE, f, out).astype(F32)


# --- R8: strip reductions feeding the validation plots ------------------------------------------

def strip_area(ncol):
    """demo_2d_LBM_JRL_C_cylinder.py:176-180: pi*((k+1)^2 - k^2), k = 0..ncol-1 (fp32)."""
    a = F32(math.pi) * np.array([i * i for i in range(ncol + 1)], dtype=F32)
    return a[1:] - a[:-1]


def strip_flux(C, mesh_grid, y0, ncol):
    """demo_2d_LBM_JRL_C_cylinder.py:344-348 (y0 = walks_int) and :355-359 (y0 = walks_int+1):
    sum over k<ncol of ((C[y0,k]-C[y0-1,k]) + (r[k+1]C[y0,k+1]-r[k]C[y0,k])/r[k]) * area[k]."""
    area = strip_area(ncol)
    dCz = C[0, 0, y0, 0:ncol] - C[0, 0, y0 - 1, 0:ncol]
    r = mesh_grid[0, 0, y0, 0:ncol + 1]
    rC = r * C[0, 0, y0, 0:ncol + 1]
    dCr = (rC[1:] - rC[:-1]) / r[:-1]
    return ((dCz + dCr) * area).sum(dtype=F32)


def strip_sum(A, y0, ncol):
    """sum_k A[y0,k]*area[k]: the balance record (demo_2d_LBM_JRL_C_cylinder.py:239-241), the
    45-degree tip current (demo_2d_LBM_JRL_C_45degree.py:353-376) and record_u / record_v
    (demo_2d_LBM_JRL_C_cylinder.py:372-377).  A is a [H,W] plane."""
    return (A[y0, 0:ncol] * strip_area(ncol)).sum(dtype=F32)


# --- R9 (extension, parity unpinned: absent from the reference) -------------------

def momentum_exchange(f_post, flags):
    """Force on all OBSTACLE cells by momentum exchange over halfway bounce-back links:
    F = sum over fluid cells x and directions q with x+e_q OBSTACLE of 2 * e_q * f_post[q](x)
    (fixed wall: the population leaving along q returns along opp(q) with the same value).
    `f_post` are post-collision populations; non-periodic neighbour lookup (links leaving the
    domain are ignored).  Returns (Fx, Fy) as float64."""
    B, _, H, W = f_post.shape
    fl = flags[:, 0]
    fx = fy = 0.0
    for q in range(1, 9):
        ex, ey = int(E[q, 0]), int(E[q, 1])
        ys, yd = slice(max(0, -ey), H - max(0, ey)), slice(max(0, ey), H - max(0, -ey))
        xs, xd = slice(max(0, -ex), W - max(0, ex)), slice(max(0, ex), W - max(0, -ex))
        link = (fl[:, ys, xs] != OBSTACLE) & (fl[:, yd, xd] == OBSTACLE)
        s = 2.0 * f_post[:, q, ys, xs][link].astype(F64).sum()
        fx += ex * s
        fy += ey * s
    return fx, fy
