"""Host-side mirror of the reference's solver objects (`src/LBM`), backed by libjrl_lbm.so.

Same class names, method names, keyword names, return arity and in-place side effects as the
reference (every method cites the reference method it replaces, paths
relative to the reference tree), so demo/*.py and batch_validation.py drive it unchanged.
Each method is at most ONE library launch (the reference: one custom kernel plus 10-80 torch
elementwise launches and one or two full copies of f) -- and when a demo calls them in the
reference's order, propagation -> [solve_boundary] -> macro_compute(_C) -> collision, the first
three only record the call and collision() launches the whole step as ONE fused kernel
(jrl_lazy.py; bit-identical to the four launches).  The explicit fused-step API and the multi-GPU
row-slab driver live in jrl_step.py / jrl_slab.py.

No CPU path: CUDA fp32 tensors only (uint8 flags); anything else raises RuntimeError.
"""
from __future__ import annotations

import ctypes
import math
import weakref

import torch

import jrl_lazy as L
import jrl_native as N
from jrl_lazy import resolve

_TWO_LAUNCHES = "axisymmetric fluid source needs the neighbours' post-stream moments: fused first half + collision"
CELL_OBSTACLE = 2
CELL_INFLOW_2 = 32
_KBC_CODES = (0, 0x85, 0x86, 0x89, 0x8A)


def _dims(f, name="f"):
    if not isinstance(f, torch.Tensor) or f.dim() != 4:
        raise RuntimeError(f"{name}: expected a [B,C,H,W] tensor")
    B, _, H, W = f.shape
    return int(B), int(H), int(W)


def _opt(t):
    """The reference passes `torch.Tensor([])` for an absent optional (propagation.cu:513-517)."""
    return None if (t is None or t.numel() == 0) else t


def _full(t, shape, name, like):
    """Broadcast `t` (e.g. a [1,2,1,1] gravity term times a scalar) to the full contiguous shape."""
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(t, dtype=like.dtype, device=like.device)
    if tuple(t.shape) != tuple(shape):
        t = t.expand(shape)
    return t.contiguous()


# --------------------------------------------------------------------------------------------------
# Streaming                                      (LBM_propagation/LBM_propagation_2d.py:95-186)
# --------------------------------------------------------------------------------------------------
class LBMPropagation2d:
    rank = 2
    type_flags = ["x", "X", "y", "Y", "z", "Z"]

    def __init__(self, tau, axisymmetric_type=0, device=torch.device("cpu"), dtype=torch.float32):
        self._tau = tau
        self.axisymmetric_type = int(axisymmetric_type)
        self.device = device
        self.dtype = dtype
        self._inflow2_cache = None  # (weakref of the flags tensor, its _version) -> bool

    def _has_inflow2(self, flags):
        # LBM_propagation_2d.py:58 syncs the host every step with `(flags == INFLOW_2).any()`;
        # the answer only changes when `flags` is written, which bumps the tensor's version counter.
        # The entry is tied to the tensor OBJECT (weak reference), not to its address: the caching
        # allocator hands a freed flags buffer of one run to the next run at the same address.
        c = self._inflow2_cache
        if c is None or c[0]() is not flags or c[1] != flags._version:
            c = self._inflow2_cache = (weakref.ref(flags), flags._version, bool((flags == CELL_INFLOW_2).any()))
        return c[2]

    def propagation(self, f, flags, phi_obs=None, rho=None, vel=None, inflow_vel=None, inflow_density=None,
                    is_convection=False):
        """Pull streaming with every link rule (jrl_stream_f32) -- recorded, not launched, when the fused step may
        follow (jrl_lazy.py).  Returns f_new, or (rho, vel, f_new) under the reference's rule
        (LBM_propagation_2d.py:52-59, 89-92); rho / vel are mutated in place at INFLOW / INFLOW_2 cells
        (propagation.cu:300-362) -- at the latest when anything looks at them."""
        f, flags, phi_obs, rho, vel, inflow_vel = (resolve(t) for t in (f, flags, phi_obs, rho, vel, inflow_vel))
        B, H, W = _dims(f)
        inflow_info = inflow_density is not None or inflow_vel is not None
        if inflow_info:
            assert rho is not None
        N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f)
        if not inflow_info:
            inflow_info = self._has_inflow2(flags)
        if not (L.ENABLED and H >= 3 and W >= 3):
            f_new = self._stream(f, flags, phi_obs, rho, vel, inflow_vel, inflow_density)
            if inflow_info:
                return (torch.Tensor([]) if rho is None else rho, torch.Tensor([]) if vel is None else vel, f_new)
            return f_new
        for t, name, shape in ((f, "f", (B, 9, H, W)), (phi_obs, "phi_obs", (B, 1, H, W)), (rho, "rho", (B, 1, H, W)),
                               (vel, "vel", (B, 2, H, W)), (inflow_vel, "inflow_vel", (2,))):
            N.tensor_ptr(t, name, shape=shape, like=f, optional=name != "f")  # validate now, as the eager call would
        P = L.PendingStep(self, f, flags, _opt(phi_obs), _opt(rho), _opt(vel), _opt(inflow_vel), inflow_density)
        f_new = P.handle("f_stream", f.shape, f)
        if inflow_info:
            r = torch.Tensor([]) if rho is None else (P.handle("rho_in", rho.shape, rho) if rho.numel() else rho)
            # vel is only written at INFLOW cells when inflow_vel is given (propagation.cu:318-321): otherwise it IS
            # the caller's tensor, nothing pending
            v = torch.Tensor([]) if vel is None else (
                P.handle("vel_in", vel.shape, vel) if (vel.numel() and P.inflow_vel is not None) else vel)
            return (r, v, f_new)
        return f_new

    def _stream(self, f, flags, phi_obs, rho, vel, inflow_vel, inflow_density):
        """One launch of jrl_stream_f32 (K1, propagation.cu:22-488)."""
        B, H, W = _dims(f)
        p_flags = N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f)
        p_f = N.tensor_ptr(f, "f", shape=(B, 9, H, W))
        p_phi = N.tensor_ptr(phi_obs, "phi_obs", shape=(B, 1, H, W), like=f, optional=True)
        p_rho = N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=f, optional=True)
        p_vel = N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=f, optional=True)
        p_iv = N.tensor_ptr(inflow_vel, "inflow_vel", shape=(2,), like=f, optional=True)
        f_new = torch.empty_like(f)
        N.call("jrl_stream_f32", p_flags, p_f, p_phi, ctypes.c_void_p(f_new.data_ptr()), p_rho, p_vel, p_iv,
               -1.0 if inflow_density is None else float(inflow_density), self.axisymmetric_type, B, H, W, like=f)
        return f_new

    def _inflow_apply(self, flags, rho, vel, inflow_vel, inflow_density):
        """Streaming's in-place writes to rho / vel alone (jrl_inflow_apply_f32)."""
        like = rho if rho is not None else vel
        B, H, W = _dims(like, "rho")
        N.call("jrl_inflow_apply_f32", N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=like),
               N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=like, optional=True),
               N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=like, optional=True),
               N.tensor_ptr(inflow_vel, "inflow_vel", shape=(2,), like=like, optional=True),
               -1.0 if inflow_density is None else float(inflow_density), B, H, W, like=like)

    def rebounce_obstacle(self, f, flags):
        """OBSTACLE cells get their populations reversed (LBM_propagation_2d.py:181-186)."""
        f, flags = resolve(f), resolve(flags)
        B, H, W = _dims(f)
        f_new = torch.empty_like(f)
        N.call("jrl_rebounce_f32", N.tensor_ptr(f, "f", shape=(B, 9, H, W)),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f),
               ctypes.c_void_p(f_new.data_ptr()), B, H, W, like=f)
        return f_new


# --------------------------------------------------------------------------------------------------
# Moments                                        (LBM_macro_compute/LBM_macro_compute_2d.py:105-361)
# --------------------------------------------------------------------------------------------------
class LBMMacroCompute2d:
    rank = 2
    type_flags = ["x", "X", "y", "Y", "z", "Z"]

    def __init__(self, Q=9, tau=1.0, density_liquid=0.265, density_gas=0.038, rho_liquid=0.265, rho_gas=0.038,
                 axisymmetric_type=0, contact_angle=0.5 * math.pi, device=torch.device("cpu"), dtype=torch.float32):
        if Q != 9:
            raise RuntimeError("D2Q9 only")
        self._Q = Q
        self._tau = tau
        self._density_liquid = density_liquid
        self._density_gas = density_gas
        self._rho_liquid = rho_liquid
        self._rho_gas = rho_gas
        self.axisymmetric_type = int(axisymmetric_type)
        self._contact_angle = contact_angle
        self.device = device
        self.dtype = dtype
        self._contact_tan_cache = None

    def _contact_tan(self):
        """tan(pi/2 - contact_angle) as the reference evaluates it (LBM_macro_compute_2d.py:246):
        fp32 tensor arithmetic for a tensor angle.  Evaluated once (one host read), not per step."""
        ca = self._contact_angle
        key = id(ca)
        if self._contact_tan_cache is None or self._contact_tan_cache[0] != key:
            if isinstance(ca, torch.Tensor):
                t = torch.tan(math.pi / 2.0 - ca.detach().to("cpu", torch.float32).reshape(-1)[0])
                val = float(t)
            else:
                val = math.tan(math.pi / 2.0 - float(ca))
            self._contact_tan_cache = (key, val)
        return self._contact_tan_cache[1]

    def macro_compute(self, dx, dt, f, rho, vel, flags, g=None, pressure=None, density=None):
        """rho = sum f, u = sum f e c / rho; OBSTACLE cells keep the old values, with the
        contact-angle wall density when contact_angle is not None (LBM_macro_compute_2d.py:316-351,
        237-271).  One launch of jrl_moments_f32 (+ a ring kernel for the wall density) -- recorded, not launched,
        when `f` is the still-unlaunched result of propagation() [+ solve_boundary()] on the same flags."""
        P = f.pending if isinstance(f, L.LazyField) else None
        if (P is not None and density is None and P.moments is None and P.owns(f, P.f_role()) and flags is P.flags
                and (P.rho_in is None or P.is_input(rho, P.rho_in, "rho_in"))
                and (P.vel_in is None or P.is_input(vel, P.vel_in, "vel_in"))):
            B, H, W = _dims(f)
            rho_t = P.rho_in if P.rho_in is not None else resolve(rho)
            vel_t = P.vel_in if P.vel_in is not None else resolve(vel)
            N.tensor_ptr(rho_t, "rho", shape=(B, 1, H, W), like=P.f_in)
            N.tensor_ptr(vel_t, "vel", shape=(B, 2, H, W), like=P.f_in)
            return P.add_moments(self, "fluid", dx, dt, rho_t, vel_t)
        f, rho, vel, flags = resolve(f), resolve(rho), resolve(vel), resolve(flags)
        rho_new, vel_new = self._moments(dx, dt, f, rho, vel, flags)
        if density is not None:
            # multiphase density remap (LBM_macro_compute_2d.py:338-349); no demo takes this branch
            density = self._density_gas + (self._density_liquid - self._density_gas) * (
                (rho_new - self._rho_gas) / (self._rho_liquid - self._rho_gas))
            return [rho_new, vel_new, density]
        return [rho_new, vel_new]

    def _moments(self, dx, dt, f, rho, vel, flags):
        B, H, W = _dims(f)
        rho_new = torch.empty((B, 1, H, W), dtype=f.dtype, device=f.device)
        vel_new = torch.empty((B, 2, H, W), dtype=f.dtype, device=f.device)
        use_contact = self._contact_angle is not None
        N.call("jrl_moments_f32", N.tensor_ptr(f, "f", shape=(B, 9, H, W)),
               N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=f), N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=f),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f),
               ctypes.c_void_p(rho_new.data_ptr()), ctypes.c_void_p(vel_new.data_ptr()),
               int(use_contact), self._contact_tan() if use_contact else 0.0, float(dx) / float(dt), B, H, W, like=f)
        return rho_new, vel_new

    def macro_compute_C(self, dx, dt, h, C, flags, vel, mesh_grid):
        """C = sum h; OBSTACLE cells keep C (LBM_macro_compute_2d.py:273-295).  Recorded, not launched, when `h` is the
        still-unlaunched result of propagation() on the same flags with `C` as its rho."""
        P = h.pending if isinstance(h, L.LazyField) else None
        if (P is not None and P.moments is None and P.wall is None and P.owns(h, "f_stream") and flags is P.flags
                and P.rho_in is not None and P.is_input(C, P.rho_in, "rho_in")):
            return P.add_moments(self, "C", dx, dt, P.rho_in, resolve(vel) if not P.is_input(vel, P.vel_in, "vel_in")
                                 else P.vel_in, mesh_grid=mesh_grid)
        return self._moments_C(resolve(h), resolve(C), resolve(flags))

    def _moments_C(self, h, C, flags):
        B, H, W = _dims(h, "h")
        C_new = torch.empty((B, 1, H, W), dtype=h.dtype, device=h.device)
        N.call("jrl_moments_C_f32", N.tensor_ptr(h, "h", shape=(B, 9, H, W)),
               N.tensor_ptr(C, "C", shape=(B, 1, H, W), like=h),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=h),
               ctypes.c_void_p(C_new.data_ptr()), 0, B, H, W, like=h)
        return C_new

    def ApplyCompute_C(self, dx, dt, h, C, flags, vel, mesh_grid):
        """The reference's CUDA twin with the axisymmetric division (macro_compute.cu:22-78)."""
        h, C, flags, vel, mesh_grid = (resolve(t) for t in (h, C, flags, vel, mesh_grid))
        B, H, W = _dims(h, "h")
        C_new = torch.empty((B, 1, H, W), dtype=h.dtype, device=h.device)
        N.call("jrl_moments_C_kernel_f32", float(dt), N.tensor_ptr(h, "h", shape=(B, 9, H, W)),
               N.tensor_ptr(C, "C", shape=(B, 1, H, W), like=h),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=h),
               N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=h, optional=True),
               N.tensor_ptr(mesh_grid, "mesh_grid", shape=(B, 2, H, W), like=h, optional=True),
               ctypes.c_void_p(C_new.data_ptr()), self.axisymmetric_type, B, H, W, like=h)
        return C_new

    def ApplyCompute_fluid(self, dx, dt, f, rho, vel, flags, g=None, pressure=None, density=None):
        """Moments without the wall-density step (macro_compute.cu:80-157)."""
        f, rho, vel, flags = resolve(f), resolve(rho), resolve(vel), resolve(flags)
        B, H, W = _dims(f)
        rho_new = torch.empty((B, 1, H, W), dtype=f.dtype, device=f.device)
        vel_new = torch.empty((B, 2, H, W), dtype=f.dtype, device=f.device)
        N.call("jrl_moments_kernel_f32", N.tensor_ptr(f, "f", shape=(B, 9, H, W)),
               N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=f), N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=f),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f),
               ctypes.c_void_p(rho_new.data_ptr()), ctypes.c_void_p(vel_new.data_ptr()),
               float(dx) / float(dt), B, H, W, like=f)
        return [rho_new, vel_new]

    def get_vort(self, vel, dx):
        """Central-difference vorticity for the PNG dumps (LBM_macro_compute_2d.py:353-361); runs
        every 10-1000 steps, off the per-step path, so it stays a handful of torch ops."""
        vel = resolve(vel)
        v = ((vel[..., 0:1, 2:, 1:-1] - vel[..., 0:1, :-2, 1:-1])
             - (vel[..., 1:2, 1:-1, 2:] - vel[..., 1:2, 1:-1, :-2])) / (2.0 * dx)
        return torch.nn.functional.pad(v, pad=(1, 1, 1, 1), mode="replicate")


# --------------------------------------------------------------------------------------------------
# Collision                      (LBM_collision/LBM_collision_2d.py, _MRT_2d.py, _KBC_2d.py)
# --------------------------------------------------------------------------------------------------
_MRT_M = [
    [1, 1, 1, 1, 1, 1, 1, 1, 1],
    [-4, -1, -1, -1, -1, 2, 2, 2, 2],
    [4, -2, -2, -2, -2, 1, 1, 1, 1],
    [0, 1, 0, -1, 0, 1, -1, -1, 1],
    [0, -2, 0, 2, 0, 1, -1, -1, 1],
    [0, 0, 1, 0, -1, 1, 1, -1, -1],
    [0, 0, -2, 0, 2, 1, 1, -1, -1],
    [0, 1, -1, 1, -1, 0, 0, 0, 0],
    [0, 0, 0, 0, 0, 1, -1, 1, -1],
]  # Lallemand-Luo moment basis in this lattice's direction order, LBM_collision_MRT_2d.py:41-57


def mrt_operator(tau, device="cpu"):
    """M^-1 S M as a host fp32 [9,9] tensor, S = diag(0, s7, s7, 0, s4, 0, s4, s7, s7), s7 = 1/tau,
    s4 = 8(2 - s7)/(8 - s7)  (LBM_collision_MRT_2d.py:36-69).  Built once per preset_KBC with the reference's own
    torch calls ON THE RUN'S DEVICE -- `torch.linalg.inv` and the 9x9 product give different last bits on CUDA
    (cuSOLVER / cuBLAS) and on the CPU (LAPACK), and the kernels multiply by exactly this matrix -- then copied to
    the host once (the C ABI takes the 81 coefficients by host pointer)."""
    M = torch.Tensor(_MRT_M).to(torch.float32).to(device).reshape(1, 1, 1, 9, 9)
    s7 = 1.0 / tau
    s4 = 8.0 * (2 - s7) / (8 - s7)
    S = torch.Tensor([0.0, s7, s7, 0.0, s4, 0.0, s4, s7, s7]).to(torch.float32).to(device).reshape(1, 1, 1, 9, 1)
    return (torch.linalg.inv(M) @ (S * M)).reshape(9, 9).to("cpu").contiguous()


class LBMCollision2d:
    """BGK (and the advection-diffusion form used for the concentration field)."""
    rank = 2
    type_flags = ["x", "X", "y", "Y", "z", "Z"]
    _op = N.OP_BGK

    def __init__(self, Q=9, tau=1.0, density_liquid=0.265, density_gas=0.038, rho_liquid=0.265, rho_gas=0.038,
                 gravity_strength=0.0, kappa=0.08, tau_f=0.7, tau_g=0.7, axisymmetric_type=0,
                 contact_angle=math.pi / 2.0, device=torch.device("cpu"), dtype=torch.float32):
        if Q != 9:
            raise RuntimeError("D2Q9 only")
        self._Q = Q
        self._tau = tau
        self._density_liquid = density_liquid
        self._density_gas = density_gas
        self._rho_liquid = rho_liquid
        self._rho_gas = rho_gas
        self._kappa = kappa
        self._tau_f = tau_f
        self._tau_g = tau_g
        self.axisymmetric_type = int(axisymmetric_type)
        self._contact_angle = contact_angle
        self.device = device
        self.dtype = dtype
        # LBM_collision_2d.py:166-171; demos form `force = density * collision._gravity`
        self._gravity = torch.tensor([0.0, -gravity_strength], dtype=dtype).reshape(1, 2, 1, 1).to(device)
        self._mrt = None

    # -- configuration ---------------------------------------------------------------------------
    def preset_KBC(self, dx, dt, tau, tau_D=None):
        """LBM_collision_2d.py:212-215."""
        self._tau = tau
        if tau_D is not None:
            self._tau_D = tau_D

    # -- single-launch operators -------------------------------------------------------------------
    def get_feq_(self, dx, dt, rho, vel, tau=None, force=None, is_convection=False):
        """Second-order equilibrium with the half-force velocity shift (LBM_collision_2d.py:236-286)."""
        rho, vel, force = resolve(rho), resolve(vel), resolve(force)
        B, H, W = _dims(vel, "vel")
        rho = _full(rho, (B, 1, H, W), "rho", vel)
        force = _full(force, (B, 2, H, W), "force", vel)
        feq = torch.empty((B, 9, H, W), dtype=vel.dtype, device=vel.device)
        N.call("jrl_feq_f32", N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=vel),
               N.tensor_ptr(vel, "vel", shape=(B, 2, H, W)),
               N.tensor_ptr(force, "force", shape=(B, 2, H, W), like=vel, optional=True),
               ctypes.c_void_p(feq.data_ptr()), float(dx), float(dt), B, H, W, like=vel)
        return feq

    def get_grad(self, rho, dx, flags):
        """9-point isotropic gradient, collision_2d.cu:22-148 via LBM_collision_2d.py:308-309."""
        rho, flags = resolve(rho), resolve(flags)
        B, H, W = _dims(rho, "rho")
        rho = rho.contiguous()
        out = torch.empty((B, 2, H, W), dtype=rho.dtype, device=rho.device)
        N.call("jrl_grad_f32", N.tensor_ptr(rho, "rho", shape=(B, 1, H, W)),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=rho),
               ctypes.c_void_p(out.data_ptr()), float(dx), self.axisymmetric_type, B, H, W, like=rho)
        return out

    def get_div(self, vel, dx, flags):
        """9-point isotropic divergence, collision_2d.cu:150-220 via LBM_collision_2d.py:311-312."""
        vel, flags = resolve(vel), resolve(flags)
        B, H, W = _dims(vel, "vel")
        vel = vel.contiguous()
        out = torch.empty((B, 1, H, W), dtype=vel.dtype, device=vel.device)
        N.call("jrl_div_f32", N.tensor_ptr(vel, "vel", shape=(B, 2, H, W)),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=vel),
               ctypes.c_void_p(out.data_ptr()), float(dx), self.axisymmetric_type, B, H, W, like=vel)
        return out

    def _operator(self, is_convection, KBC_type):
        return self._op, 0, (self._tau_D if is_convection else self._tau)

    def collision(self, dx, dt, f, rho, vel, flags, force=None, mesh_grid=None, is_convection=False, KBC_type=None,
                  **unused):
        """f' = f + Omega(f, feq) + dt G; OBSTACLE cells keep f (LBMCollision2d.collision, LBM_collision_2d.py:473-555,
        including get_feq_ and the axisymmetric source :343-471 with its gradient / divergence stencils).
        When f, rho, vel are the still-unlaunched results of this step's propagation() [+ solve_boundary()] +
        macro_compute(_C)(), the WHOLE step is launched here as one fused kernel (jrl_step_f32, 72 B per cell instead of
        four passes); otherwise whatever is pending is launched call by call and this is one launch of jrl_collide_f32."""
        op, kbc, tau = self._operator(is_convection, KBC_type)
        P = f.pending if isinstance(f, L.LazyField) else None
        if P is not None:
            why = self._fusable(P, dx, dt, f, rho, vel, flags, mesh_grid, is_convection, tau)
            if why is None:
                return self._fused_step(P, dx, dt, op, kbc, tau, force, mesh_grid, is_convection)
            if why is _TWO_LAUNCHES:
                # axisymmetric FLUID form (JRL stage A): its source term takes gradients of the NEW moments, so the
                # step is two launches -- streaming + wall + moments fused (JRL_OP_NONE), then this collision
                f = self._fused_step(P, dx, dt, N.OP_NONE, 0, tau, None, None, False)
                rho, vel = P.real["rho"], P.real["vel"]
            else:
                L._why(why)
        f, rho, vel, flags, force, mesh_grid = (resolve(t) for t in (f, rho, vel, flags, force, mesh_grid))
        return self._collide(dx, dt, f, rho, vel, flags, force, mesh_grid, is_convection, op, kbc, tau)

    def _collide(self, dx, dt, f, rho, vel, flags, force, mesh_grid, is_convection, op, kbc, tau, slab=None, f_out=None):
        """One launch of jrl_collide_f32 on ordinary tensors.  slab = (row0, nrows, H_global): the tensors are row slabs
        [.., nrows + 2, W] with exchanged ghost rows (jrl_slab.AxisymFluidSlab); only the owned rows of f_out are written."""
        B, H, W = _dims(f)
        force = _full(force, (B, 2, H, W), "force", f)
        d = N.CollideDesc()
        d.B, d.H, d.W = B, H, W
        if slab is not None:
            d.slab_row0, d.slab_rows, d.H = int(slab[0]), int(slab[1]), int(slab[2])
        d.axisym = self.axisymmetric_type
        d.op, d.kbc_type, d.is_convection = op, kbc, int(bool(is_convection))
        d.dx, d.dt, d.tau = float(dx), float(dt), float(tau)
        d.mrt = self._mrt.data_ptr() if (op == N.OP_MRT and self._mrt is not None) else None
        d.f = N.tensor_ptr(f, "f", shape=(B, 9, H, W))
        d.rho = N.tensor_ptr(rho, "rho", shape=(B, 1, H, W), like=f)
        d.vel = N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=f)
        d.flags = N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f)
        d.force = N.tensor_ptr(force, "force", shape=(B, 2, H, W), like=f, optional=True)
        d.mesh_grid = N.tensor_ptr(mesh_grid, "mesh_grid", shape=(B, 2, H, W), like=f, optional=True)
        if f_out is None:
            f_out = torch.empty_like(f)
        d.f_out = N.tensor_ptr(f_out, "f_out", shape=(B, 9, H, W), like=f)
        N.call("jrl_collide_f32", ctypes.byref(d), like=f)
        return f_out

    # -- the fused step behind the four calls (jrl_lazy.py) --------------------------------------------------------
    def _fusable(self, P, dx, dt, f, rho, vel, flags, mesh_grid, is_convection, tau):
        """None when (f, rho, vel) are exactly this step's recorded chain and jrl_step_f32 computes it; else the reason
        the calls are replayed one launch each."""
        m = P.moments
        if P.launched or m is None or not P.owns(f, P.f_role()) or flags is not P.flags:
            return "collision() input is not this step's propagation -> macro_compute chain"
        B, H, W = _dims(P.f_in)
        if float(m["dx"]) != float(dx) or float(m["dt"]) != float(dt):
            return "dx / dt differ between the calls"
        if is_convection:
            if m["kind"] != "C" or not P.owns(rho, "C_clamp" if P.clamp else "C"):
                return "concentration collision without this step's macro_compute_C result"
            if P.vel_in is None or not P.is_input(vel, P.vel_in, "vel_in") or m["vel"] is not P.vel_in:
                return "advecting velocity differs between propagation and collision"
            if self.axisymmetric_type != N.AXI_NOT and mesh_grid is None:
                return "axisymmetric concentration step without mesh_grid"
        else:
            if m["kind"] != "fluid" or not P.owns(rho, "rho") or not P.owns(vel, "vel"):
                return "collision without this step's macro_compute result"
            if (P.rho_in is None or P.vel_in is None) and self._has_inflow(P.flags):
                return "INFLOW cells with rho / vel not given to propagation()"
            if m["obj"]._contact_angle is not None and (H < 4 or W < 4):
                return "grid too small for the fused contact-angle ring"
        w = P.wall
        if w is not None:
            if is_convection or float(w["tau"]) != float(tau) or float(w["dx"]) != float(dx) or float(w["dt"]) != float(dt):
                return "moving-wall parameters differ from the collision's"
            if w["vel"] is not m["vel"]:
                return "moving wall and macro_compute were given different previous velocities"
        if self.axisymmetric_type != P.prop.axisymmetric_type or self.axisymmetric_type != m["obj"].axisymmetric_type:
            return "solver objects disagree on the symmetry axis"
        if not is_convection and self.axisymmetric_type != N.AXI_NOT:
            return _TWO_LAUNCHES
        return None

    def _has_inflow(self, flags):
        c = getattr(self, "_inflow_cache", None)
        if c is None or c[0]() is not flags or c[1] != flags._version:
            c = self._inflow_cache = (weakref.ref(flags), flags._version,
                                      bool(((flags == 8) | (flags == CELL_INFLOW_2)).any()))
        return c[2]

    def _fused_step(self, P, dx, dt, op, kbc, tau, force, mesh_grid, is_convection):
        import jrl_step
        from types import SimpleNamespace
        m = P.moments
        ca = None if is_convection else m["obj"]._contact_angle
        key = (op, kbc, float(tau), bool(is_convection), float(dx), float(dt), id(ca), self.axisymmetric_type)
        cache = self.__dict__.setdefault("_fused_cache", {})
        fs = cache.get(key)
        if fs is None:
            par = SimpleNamespace(tau=tau, axisymmetric_type=self.axisymmetric_type, contact_angle=ca, device=P.f_in.device)
            name = {N.OP_BGK: "bgk", N.OP_MRT: "mrt", N.OP_KBC: "kbc", N.OP_NONE: "none"}[op]
            fs = jrl_step.FusedStep(par, operator=name, kbc_type=kbc if op == N.OP_KBC else None, tau=tau,
                                    tau_D=tau if is_convection else None, is_convection=is_convection, dx=dx, dt=dt,
                                    contact_angle=ca, mrt=self._mrt if op == N.OP_MRT else None)
            if len(cache) >= 8:
                cache.clear()
            cache[key] = fs
        P.check_unmodified()
        w = P.wall
        out = fs.step(P.f_in, P.flags, m["rho"], m["vel"], phi_obs=P.phi_obs, force=resolve(force),
                      mesh_grid=resolve(mesh_grid), inflow_vel=P.inflow_vel, inflow_density=P.inflow_density,
                      use_wall=w is not None, wall_phi=None if w is None else w["phi_obs"],
                      obs_vel=None if w is None else w["obs_vel"], clamp01=P.clamp)
        if is_convection:
            P.publish_fused(out[1], None)
        else:
            P.publish_fused(out[1], out[2])
        return out[0]


class LBMCollisionMRT2d(LBMCollision2d):
    """Multiple-relaxation-time operator (LBM_collision_MRT_2d.py:11-184); what demo_2d_LBM_KBC.py
    and demo_2d_LBM_cylinder.py really run (their KBC_type argument is ignored)."""
    _op = N.OP_MRT

    def preset_KBC(self, dx, dt, tau, tau_D=None):
        self._tau = tau
        self._mrt = mrt_operator(tau, self.device)
        self.Minv_S_M = self._mrt.reshape(1, 1, 1, 9, 9).to(self.device)
        if tau_D is not None:
            self._tau_D = tau_D

    def _operator(self, is_convection, KBC_type):
        assert not is_convection  # LBM_collision_MRT_2d.py:138
        if self._mrt is None:
            raise AttributeError("LBMCollisionMRT2d: call preset_KBC(dx, dt, tau) first")  # reference: no Minv_S_M yet
        return N.OP_MRT, 0, self._tau


class LBMCollisionKBC2d(LBMCollision2d):
    """Entropic KBC stabiliser (LBM_collision_KBC_2d.py:11-317)."""
    _op = N.OP_KBC

    def _operator(self, is_convection, KBC_type):
        assert not is_convection
        code = 0 if KBC_type is None else int(KBC_type)
        if code not in _KBC_CODES:
            raise RuntimeError("Please specific a right KBC Type value")  # LBM_collision_KBC_2d.py:235
        return N.OP_KBC, code, self._tau


# --------------------------------------------------------------------------------------------------
# Moving-wall coupling            (LBM_solid_fluid_coupling/LBM_boundary_2d.py:76-158)
# --------------------------------------------------------------------------------------------------
class LBMBoundary2d:
    rank = 2
    type_flags = ["x", "X", "y", "Y", "z", "Z"]

    def __init__(self, tau, axisymmetric_type=0, device=torch.device("cpu"), dtype=torch.float32):
        self._tau = tau
        self.axisymmetric_type = int(axisymmetric_type)
        self.device = device
        self.dtype = dtype

    def solve_boundary(self, dt, dx, rho, vel, flags, f, phi_obs, obs_vel, tau=None, is_convection=False):
        """Regularised moving-wall reconstruction at fluid cells next to obstacles; one launch of
        jrl_wall_f32 (solid_fluid_coupling.cu:25-252) -- recorded, not launched, when `f` is the still-unlaunched
        result of propagation() on the same flags.  `rho` and `is_convection` are unused by the reference kernel too."""
        tau = self._tau if tau is None else tau
        P = f.pending if isinstance(f, L.LazyField) else None
        phi_obs, obs_vel = resolve(phi_obs), _opt(resolve(obs_vel))
        if (P is not None and P.wall is None and P.moments is None and P.owns(f, "f_stream") and flags is P.flags
                and (P.vel_in is None or P.is_input(vel, P.vel_in, "vel_in"))):
            B, H, W = _dims(f)
            vel_t = P.vel_in if P.vel_in is not None else resolve(vel)
            N.tensor_ptr(vel_t, "vel", shape=(B, 2, H, W), like=P.f_in)
            N.tensor_ptr(phi_obs, "phi_obs", shape=(B, 1, H, W), like=P.f_in)
            N.tensor_ptr(obs_vel, "obs_vel", shape=(B, 2, H, W), like=P.f_in, optional=True)
            return P.add_wall(self, dt, dx, tau, rho, vel_t, phi_obs, obs_vel)
        return self._wall(dt, dx, tau, resolve(vel), resolve(flags), resolve(f), phi_obs, obs_vel)

    def _wall(self, dt, dx, tau, vel, flags, f, phi_obs, obs_vel):
        B, H, W = _dims(f)
        f_new = torch.empty_like(f)
        N.call("jrl_wall_f32", float(dt), float(dx), float(tau), N.tensor_ptr(vel, "vel", shape=(B, 2, H, W), like=f),
               N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f),
               N.tensor_ptr(f, "f", shape=(B, 9, H, W)), ctypes.c_void_p(f_new.data_ptr()),
               N.tensor_ptr(phi_obs, "phi_obs", shape=(B, 1, H, W), like=f),
               N.tensor_ptr(obs_vel, "obs_vel", shape=(B, 2, H, W), like=f, optional=True),
               self.axisymmetric_type, B, H, W, like=f)
        return f_new


# --------------------------------------------------------------------------------------------------
# R8 / R9 reductions (new entry points; the demos do these with ~20 tiny torch launches per step)
# --------------------------------------------------------------------------------------------------
def strip_flux(C, mesh_grid, y0, ncol, out=None):
    """[flux, plain] of demo_2d_LBM_JRL_C_cylinder.py:344-362 / 45degree.py:353-376 for row y0 on
    the device (one warp); returns a device [2] tensor (no host sync)."""
    B, H, W = _dims(C, "C")
    if out is None:
        out = torch.empty(2, dtype=C.dtype, device=C.device)
    N.call("jrl_strip_flux_f32", N.tensor_ptr(C, "C", shape=(B, 1, H, W)),
           N.tensor_ptr(mesh_grid, "mesh_grid", shape=(B, 2, H, W), like=C), int(y0), int(ncol),
           ctypes.c_void_p(out.data_ptr()), H, W, like=C)
    return out


def momentum_exchange(f_post, flags):
    """Extension (absent from the reference): momentum-exchange force on all OBSTACLE
    cells, device [2] float64 (Fx, Fy)."""
    B, H, W = _dims(f_post)
    out = torch.zeros(2, dtype=torch.float64, device=f_post.device)
    N.call("jrl_momentum_exchange_f32", N.tensor_ptr(f_post, "f", shape=(B, 9, H, W)),
           N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, H, W), like=f_post),
           ctypes.c_void_p(out.data_ptr()), B, H, W, like=f_post)
    return out


def tip_prepare_conc(mesh_grid, flags, C, vel0, vel1, t, box_center, box_radius, walks_int, Rg_radius, radius_obs_int,
                     vel_obs_conv, is_negative=True, phi_out=None, vel_out=None, shape=N.TIP_BOX):
    """Device-side tip motion for the JRL concentration walk loop (jrl_tip_prepare_conc_f32): one launch instead
    of the demo's ~15 torch ops per step (demo_2d_LBM_JRL_C_cylinder.py:283-307).  flags and C are updated in
    place; returns (phi_obs, C_vel).  shape: N.TIP_BOX (cylinder, infinite plane) or N.TIP_CONE45
    (demo_2d_LBM_JRL_C_45degree.py:31-46).  vel0 / vel1: [2,H,W] (or [1,2,H,W]) velocity snapshots and the interpolation
    weight t (vel1 None: vel0 alone); box_center: the demo's fp32 [2] tensor or a pair of floats."""
    B, H, W = _dims(C, "C")
    if B != 1:
        raise RuntimeError("tip_prepare_conc: batch size 1 only")
    d = N.TipDesc()
    d.H, d.W, d.is_negative, d.shape = H, W, int(bool(is_negative)), int(shape)
    d.walks_int, d.Rg_radius, d.radius_obs_int = int(walks_int), int(Rg_radius), int(radius_obs_int)
    bc = box_center.detach().to("cpu", torch.float32).tolist() if isinstance(box_center, torch.Tensor) else list(box_center)
    d.box_cx, d.box_cy, d.box_radius = float(bc[0]), float(bc[1]), float(box_radius)
    d.vel_obs_conv = float(vel_obs_conv)
    d.w0, d.w1 = float(1.0 - t), float(t)
    d.mesh_grid = N.tensor_ptr(mesh_grid, "mesh_grid", shape=(1, 2, H, W), like=C)
    d.flags = N.tensor_ptr(flags, "flags", torch.uint8, (1, 1, H, W), like=C)
    d.C = N.tensor_ptr(C, "C", shape=(1, 1, H, W))
    v0 = vel0.reshape(1, 2, H, W)
    d.vel0 = N.tensor_ptr(v0, "vel0", shape=(1, 2, H, W), like=C)
    v1 = None if vel1 is None else vel1.reshape(1, 2, H, W)
    d.vel1 = N.tensor_ptr(v1, "vel1", shape=(1, 2, H, W), like=C, optional=True)
    if phi_out is None:
        phi_out = torch.empty_like(C)
    if vel_out is None:
        vel_out = torch.empty((1, 2, H, W), dtype=C.dtype, device=C.device)
    d.phi_obs = N.tensor_ptr(phi_out, "phi_obs", shape=(1, 1, H, W), like=C)
    d.C_vel = N.tensor_ptr(vel_out, "C_vel", shape=(1, 2, H, W), like=C)
    N.call("jrl_tip_prepare_conc_f32", ctypes.byref(d), like=C)
    # the kernel wrote flags behind torch's back: bump the version counter so that cached tile masks are rebuilt
    try:
        torch.autograd.graph.increment_version(flags)
    except AttributeError:
        flags.add_(0)
    return phi_out, vel_out
