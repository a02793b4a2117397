"""The fused lattice update behind one call: jrl_step_f32 (include/jrl_lbm.h).

One launch = prop.propagation -> [coupling.solve_boundary] -> macro.macro_compute(_C) ->
collision.collision of the reference's demo loops (demo_2d_LBM.py:101-119,
demo_2d_LBM_cylinder.py:138-172, demo_2d_LBM_JRL_C_cylinder.py:313-326), bit-identical to
calling the four solver objects in that order, at 72 B of population traffic per lattice
update instead of four full passes.  The axisymmetric FLUID collision (JRL stage A) needs
neighbours' post-stream moments and is the one combination that stays on the solver-object
path (the library rejects it).
"""
from __future__ import annotations

import ctypes
import math
import weakref

import torch

import jrl_native as N
from jrl_solver import _KBC_CODES, _dims, _full, mrt_operator

_OPS = {"bgk": N.OP_BGK, "mrt": N.OP_MRT, "kbc": N.OP_KBC, "none": N.OP_NONE}


class FusedStep:
    """Configured once per run (operator, tau, symmetry, contact angle); `step()` per time step.

    parameters : SimulationParameters (tau, axisymmetric_type, contact_angle are read)
    operator   : 'bgk' | 'mrt' | 'kbc' | 'none' (streaming + boundaries + moments, no collision: the first half of the
                 axisymmetric fluid step, whose collision needs the neighbours' new moments and is a second launch)
    is_convection : advection-diffusion form for the concentration field (BGK, relaxation tau_D)
    """

    def __init__(self, parameters, operator="bgk", kbc_type=None, tau=None, tau_D=None, is_convection=False,
                 dx=1.0, dt=1.0, contact_angle="from_parameters", mrt=None):
        self.axisymmetric_type = int(parameters.axisymmetric_type)
        self.op = _OPS[operator]
        self.kbc_type = 0 if kbc_type is None else int(kbc_type)
        if self.op == N.OP_KBC and self.kbc_type not in _KBC_CODES:
            raise RuntimeError("Please specific a right KBC Type value")
        self.is_convection = bool(is_convection)
        if self.is_convection and self.op != N.OP_BGK:
            raise RuntimeError("the concentration step is BGK only (LBM_collision_MRT_2d.py:138)")
        self.tau = float(parameters.tau if tau is None else tau)
        if self.is_convection:
            if tau_D is None:
                raise RuntimeError("is_convection needs tau_D")
            self.tau = float(tau_D)
        self.dx, self.dt = float(dx), float(dt)
        self._mrt = (mrt if mrt is not None else mrt_operator(self.tau, parameters.device)) if self.op == N.OP_MRT else None
        ca = parameters.contact_angle if isinstance(contact_angle, str) else contact_angle
        self.use_contact = (ca is not None) and not self.is_convection
        if self.use_contact:
            if isinstance(ca, torch.Tensor):
                self.contact_tan = float(torch.tan(math.pi / 2.0 - ca.detach().to("cpu", torch.float32).reshape(-1)[0]))
            else:
                self.contact_tan = math.tan(math.pi / 2.0 - float(ca))
        else:
            self.contact_tan = 0.0
        self.use_tile_mask = True
        self._worklists = {}
        self._masks = {}  # id(flags) -> (weakref(flags), flags._version, mask tensor)

    def invalidate_masks(self):
        """Forget the cached tile masks (needed only if a flags tensor was rewritten behind torch's back,
        i.e. without bumping its version counter)."""
        self._masks.clear()

    def _tile_mask(self, flags, B, Hs, W):
        """Packed all-FLUID summary of `flags` for the pipelined kernel (jrl_tile_mask_u8), recomputed
        only when the tensor was written since the last call (torch's version counter)."""
        # The entry belongs to the tensor OBJECT (weak reference), not to its address: the caching allocator gives a
        # freed flags buffer of one run to the next run at the same address with the same version count, and a stale
        # all-FLUID mask would make the pipelined kernel skip flag rows that now hold obstacles.
        key = id(flags)
        hit = self._masks.get(key)
        if hit is not None and hit[0]() is flags and hit[1] == flags._version:
            return hit[2]
        n = N.load().jrl_tile_mask_bytes(B, Hs, W)
        if n <= 0:
            return None
        mask = (hit[2] if hit is not None and hit[0]() is flags and hit[2].numel() == n
                else torch.empty((n,), dtype=torch.uint8, device=flags.device))
        N.call("jrl_tile_mask_u8", ctypes.c_void_p(flags.data_ptr()), ctypes.c_void_p(mask.data_ptr()), B, Hs, W, like=flags)
        self._masks = {k: v for k, v in self._masks.items() if v[0]() is not None}  # drop entries of dead tensors
        if len(self._masks) >= 16:
            self._masks.clear()
        self._masks[key] = (weakref.ref(flags), flags._version, mask)
        return mask

    def step(self, f, flags, rho=None, vel=None, *, f_out=None, rho_out=None, vel_out=None, store_moments=True,
             phi_obs=None, force=None, mesh_grid=None, inflow_vel=None, inflow_density=None, use_wall=False,
             obs_vel=None, wall_phi=None, clamp01=False, slab=None, rows=None, force_vec=0, sums=None):
        """One lattice update f -> f_out.  Returns (f_out, rho_out, vel_out) for the fluid form and
        (h_out, C_out) for the concentration form (`rho` is C, `vel` the advecting velocity).
        Outputs are allocated when not given (pass your own ping-pong buffers to avoid that);
        store_moments=False skips the rho/vel stores (pure 72 B/LUP path).
        slab = (row0, nrows, H): tensors are [B,C,nrows+2,W] row slabs of a global H-row grid, with
        ghost rows (multi-GPU); wall_phi: phi_obs for the wall step when streaming runs without it;
        rows = (begin, end): update only these global rows;
        sums: float64 [SUM_SLOTS, 4] device tensor this launch adds {sum rho, sum rho ux, sum rho uy,
        cells} to (see new_sums / total_sums)."""
        d, f_out, rho_out, vel_out, _keep = self._desc(f, flags, rho, vel, f_out, rho_out, vel_out, store_moments, phi_obs,
                                                       force, mesh_grid, inflow_vel, inflow_density, use_wall, obs_vel,
                                                       wall_phi, clamp01, slab, rows, force_vec, sums)
        N.call("jrl_step_f32", ctypes.byref(d), like=f)
        if self.is_convection:
            return f_out, rho_out
        return f_out, rho_out, vel_out

    def _desc(self, f, flags, rho, vel, f_out, rho_out, vel_out, store_moments, phi_obs, force, mesh_grid, inflow_vel,
              inflow_density, use_wall, obs_vel, wall_phi, clamp01, slab, rows, force_vec, sums, sums_shape=None):
        """Validated jrl_step_desc for these tensors (+ the output tensors it allocated and the objects that
        must stay alive until the launch has been issued)."""
        B, Hs, W = _dims(f)
        H = Hs
        d = N.StepDesc()
        if slab is not None:
            d.slab_row0, d.slab_rows, H = int(slab[0]), int(slab[1]), int(slab[2])
            if Hs != d.slab_rows + 2:
                raise RuntimeError("slab tensors must hold nrows + 2 rows")
        if rows is not None:
            d.row_begin, d.row_end = int(rows[0]), int(rows[1])
        d.B, d.H, d.W = B, H, W
        d.axisym = self.axisymmetric_type
        d.op, d.kbc_type, d.is_convection = self.op, self.kbc_type, int(self.is_convection)
        d.use_contact, d.contact_tan = int(self.use_contact), self.contact_tan
        d.clamp01 = int(bool(clamp01))
        d.dx, d.dt, d.tau = self.dx, self.dt, self.tau
        d.mrt = self._mrt.data_ptr() if self._mrt is not None else None
        d.flags = N.tensor_ptr(flags, "flags", torch.uint8, (B, 1, Hs, W), like=f)
        d.f_in = N.tensor_ptr(f, "f", shape=(B, 9, Hs, W))
        if f_out is None:
            f_out = torch.empty_like(f)
        d.f_out = N.tensor_ptr(f_out, "f_out", shape=(B, 9, Hs, W), like=f)
        d.phi_obs = N.tensor_ptr(phi_obs, "phi_obs", shape=(B, 1, Hs, W), like=f, optional=True)
        d.rho = N.tensor_ptr(rho, "rho", shape=(B, 1, Hs, W), like=f, optional=True)
        d.vel = N.tensor_ptr(vel, "vel", shape=(B, 2, Hs, W), like=f, optional=True)
        if store_moments or self.use_contact:
            if rho_out is None:
                rho_out = torch.empty((B, 1, Hs, W), dtype=f.dtype, device=f.device)
            if vel_out is None and not self.is_convection:
                vel_out = torch.empty((B, 2, Hs, W), dtype=f.dtype, device=f.device)
        d.rho_out = N.tensor_ptr(rho_out, "rho_out", shape=(B, 1, Hs, W), like=f, optional=True)
        d.vel_out = N.tensor_ptr(vel_out, "vel_out", shape=(B, 2, Hs, W), like=f, optional=True)
        force = _full(force, (B, 2, Hs, W), "force", f)
        d.force = N.tensor_ptr(force, "force", shape=(B, 2, Hs, W), like=f, optional=True)
        d.mesh_grid = N.tensor_ptr(mesh_grid, "mesh_grid", shape=(B, 2, Hs, W), like=f, optional=True)
        d.inflow_vel = N.tensor_ptr(inflow_vel, "inflow_vel", shape=(2,), like=f, optional=True)
        d.inflow_density = -1.0 if inflow_density is None else float(inflow_density)
        d.use_wall = int(bool(use_wall))
        d.obs_vel = N.tensor_ptr(obs_vel, "obs_vel", shape=(B, 2, Hs, W), like=f, optional=True)
        d.force_vec = int(force_vec)
        d.sums = N.tensor_ptr(sums, "sums", torch.float64, sums_shape or (N.SUM_SLOTS, 4), like=f, optional=True)
        d.wall_phi = N.tensor_ptr(wall_phi, "wall_phi", shape=(B, 1, Hs, W), like=f, optional=True)
        mask = wl = None
        if (self.use_tile_mask and W % 16 == 0 and W >= 512 and flags.data_ptr() % 16 == 0
                and (force_vec == 8 or (force_vec == 0 and self.op in (N.OP_BGK, N.OP_MRT)))):
            mask = self._tile_mask(flags, B, Hs, W)
            wl = self._worklist(flags.device, B, Hs, W)
        d.tile_mask = None if mask is None else mask.data_ptr()
        if wl is not None:
            d.worklist, d.worklist_bytes = wl.data_ptr(), wl.numel() * 4
        return d, f_out, rho_out, vel_out, (force, mask, wl)

    def _worklist(self, device, B, Hs, W):
        """Device scratch through which the pipelined kernel hands its boundary groups to the second launch
        (jrl_step_desc.worklist); one buffer per device and extent, reused by every step."""
        key = (str(device), B, Hs, W)
        wl = self._worklists.get(key)
        if wl is None:
            n = int(N.load().jrl_worklist_bytes(B, Hs, W))
            wl = self._worklists[key] = torch.zeros(((n + 3) // 4,), dtype=torch.int32, device=device)
        return wl

    def run(self, f, flags, rho=None, vel=None, nsteps=1, *, f_alt=None, rho_alt=None, vel_alt=None, store_every=0,
            phi_obs=None, force=None, mesh_grid=None, inflow_vel=None, inflow_density=None, use_wall=False, obs_vel=None,
            wall_phi=None, clamp01=False, force_vec=0, sums=None):
        """`nsteps` lattice updates in ONE library call (jrl_run_f32): the demo's time loop below the C ABI, for
        grids small enough to be paced by the interpreter.  The per-step inputs (flags, phi_obs, force, ...) stay
        fixed for the whole call.  f_alt / rho_alt / vel_alt: the second buffer of each ping-pong pair (allocated
        when not given).  When `rho` / `vel` are given they are state and ping-pong with the populations; when
        None the moments are outputs stored every `store_every`-th step and at the last one.
        sums: None, a [SUM_SLOTS, 4] accumulator, or a [nsteps, SUM_SLOTS, 4] per-step series (float64).
        Returns the newest (f, rho, vel) -- (h, C) in the concentration form."""
        nsteps = int(nsteps)
        if f_alt is None:
            f_alt = torch.empty_like(f)
        series = sums is not None and sums.dim() == 3
        d, f_alt, rho_alt, vel_alt, _keep = self._desc(
            f, flags, rho, vel, f_alt, rho_alt, vel_alt, True, phi_obs, force, mesh_grid, inflow_vel, inflow_density,
            use_wall, obs_vel, wall_phi, clamp01, None, None, force_vec, sums,
            sums_shape=(nsteps, N.SUM_SLOTS, 4) if series else None)
        N.call("jrl_run_f32", ctypes.byref(d), nsteps, int(store_every), N.SUM_SLOTS * 4 if series else 0, like=f)
        N.launch_count += max(0, nsteps - 1)
        odd = nsteps % 2 == 1
        f_new = f_alt if odd else f
        state_rho = rho is not None
        state_vel = vel is not None and not self.is_convection
        rho_new = (rho_alt if odd else rho) if state_rho else rho_alt
        if self.is_convection:
            return f_new, rho_new
        vel_new = (vel_alt if odd else vel) if state_vel else vel_alt
        return f_new, rho_new, vel_new


def conc_walk(fs, h, flags, C, mesh_grid, vel_snaps, walks, nsteps, *, box_center_x, box_radius, Rg_radius, radius_obs_int,
              vel_obs_conv, inflow_height, x_interval=0.1, is_negative=True, clamp01=True, dx=1.0, dt=1.0, records=None,
              shape=N.TIP_BOX, demo_clock=None):
    """`nsteps` of the JRL concentration demos' walk loop in ONE library call (jrl_conc_walk_f32): device-side
    tip motion, fused concentration step, the two recorded strip fluxes and the tip advance, per step, with no
    interpreter in the loop (demo_2d_LBM_JRL_C_cylinder.py:281-362).  `fs`: a FusedStep built with
    is_convection=True; flags and C are updated in place / ping-ponged.  Returns
    (h, C, records [nsteps,4] = inflow flux, inflow sum, current flux, current sum, walks, phi_obs, C_vel).
    shape: N.TIP_BOX / N.TIP_CONE45.  demo_clock: None, or a dict(conv_timer, dt_conv_real, vel_obs_real,
    radius_obs_real, radius_obs) to index the velocity snapshots exactly as the demo's python does (:294-303, :342;
    no clamp -- running past the stage-A cache raises like the demo's IndexError, after the steps that did run;
    the dict is updated with the new conv_timer and `steps_done`)."""
    if not fs.is_convection:
        raise RuntimeError("conc_walk needs a FusedStep with is_convection=True")
    B, H, W = _dims(h)
    nsteps = int(nsteps)
    h_alt, C_alt = torch.empty_like(h), torch.empty_like(C)
    phi = torch.empty_like(C)
    C_vel = torch.empty((1, 2, H, W), dtype=h.dtype, device=h.device)
    if records is None:
        records = torch.zeros((max(nsteps, 1), 4), dtype=torch.float32, device=h.device)
    d, _, _, _, keep = fs._desc(h, flags, C, C_vel, h_alt, C_alt, None, True, phi, None, mesh_grid, None, 0, False, None, None,
                                clamp01, None, None, 0, None)
    w = N.WalkDesc()
    w.step = d
    t = w.tip
    t.H, t.W, t.is_negative, t.shape = H, W, int(bool(is_negative)), int(shape)
    t.Rg_radius, t.radius_obs_int = int(Rg_radius), int(radius_obs_int)
    t.box_cx, t.box_radius, t.vel_obs_conv = float(box_center_x), float(box_radius), float(vel_obs_conv)
    t.mesh_grid = d.mesh_grid
    w.walks, w.vel_obs_conv, w.dx, w.dt = float(walks), float(vel_obs_conv), float(dx), float(dt)
    w.box_radius, w.inflow_height, w.x_interval = float(box_radius), float(inflow_height), float(x_interval)
    snaps = vel_snaps.contiguous()
    w.vel_snaps = N.tensor_ptr(snaps, "vel_snaps", shape=(snaps.shape[0], 2, H, W), like=h)
    w.n_snaps = int(snaps.shape[0])
    w.records = N.tensor_ptr(records, "records", shape=(max(nsteps, 1), 4), like=h)
    if demo_clock is not None:
        w.conv_timer, w.dt_conv_real = float(demo_clock["conv_timer"]), float(demo_clock["dt_conv_real"])
        w.vel_obs_real, w.radius_obs_real = float(demo_clock["vel_obs_real"]), float(demo_clock["radius_obs_real"])
        w.radius_obs = float(demo_clock["radius_obs"])
    try:
        N.call("jrl_conc_walk_f32", ctypes.byref(w), nsteps, like=h)
    finally:
        if demo_clock is not None:
            demo_clock["conv_timer"], demo_clock["steps_done"] = float(w.conv_timer), int(w.steps_done)
    N.launch_count += max(0, 4 * nsteps - 1)
    try:
        torch.autograd.graph.increment_version(flags)
    except AttributeError:
        flags.add_(0)
    odd = nsteps % 2 == 1
    return (h_alt if odd else h), (C_alt if odd else C), records, float(w.walks), phi, C_vel


def fluid_walk(collision, f, flags, rho, vel, mesh_grid, force, walks, nsteps, *, box_center_x, box_radius, Rg_radius, vel_obs,
               tau, dx=1.0, dt=1.0, shape=N.TIP_BOX):
    """`nsteps` of the JRL fluid demos' loop (stage A, demo_2d_LBM_JRL_fluid_cylinder.py:203-264) in ONE library call
    (jrl_fluid_walk_f32): device-side tip motion, streaming with interpolated bounce-back, moving-wall reconstruction,
    moments, BGK collision with the axisymmetric source, tip advance.  `collision`: the LBMCollision2d object (axisymmetric
    type and operator are read from it).  flags and the moment tensors are updated in place / ping-ponged.
    Returns (f, rho, vel, walks, phi_obs, obs_vel)."""
    B, H, W = _dims(f)
    if B != 1:
        raise RuntimeError("fluid_walk: batch size 1 only")
    w = N.FluidWalkDesc()
    c = w.collide
    c.B, c.H, c.W = 1, H, W
    c.axisym = int(collision.axisymmetric_type)
    c.op, c.kbc_type, c.is_convection = N.OP_BGK, 0, 0
    c.dx, c.dt, c.tau = float(dx), float(dt), float(tau)
    c.flags = N.tensor_ptr(flags, "flags", torch.uint8, (1, 1, H, W), like=f)
    force = _full(force, (1, 2, H, W), "force", f)
    c.force = N.tensor_ptr(force, "force", shape=(1, 2, H, W), like=f, optional=True)
    c.mesh_grid = N.tensor_ptr(mesh_grid, "mesh_grid", shape=(1, 2, H, W), like=f)
    phi, obs_vel = torch.empty_like(rho), torch.empty_like(vel)
    t = w.tip
    t.H, t.W, t.Rg_radius, t.shape = H, W, int(Rg_radius), int(shape)
    t.box_cx, t.box_radius, t.vel_obs = float(box_center_x), float(box_radius), float(vel_obs)
    t.mesh_grid = c.mesh_grid
    t.phi_obs = N.tensor_ptr(phi, "phi_obs", shape=(1, 1, H, W), like=f)
    t.obs_vel = N.tensor_ptr(obs_vel, "obs_vel", shape=(1, 2, H, W), like=f)
    w.walks, w.vel_obs, w.box_radius = float(walks), float(vel_obs), float(box_radius)
    fb, fc = torch.empty_like(f), torch.empty_like(f)
    rho2, vel2 = torch.empty_like(rho), torch.empty_like(vel)
    w.f[0] = N.tensor_ptr(f, "f", shape=(1, 9, H, W)).value
    w.f[1], w.f[2] = fb.data_ptr(), fc.data_ptr()
    w.rho[0] = N.tensor_ptr(rho, "rho", shape=(1, 1, H, W), like=f).value
    w.rho[1] = rho2.data_ptr()
    w.vel[0] = N.tensor_ptr(vel, "vel", shape=(1, 2, H, W), like=f).value
    w.vel[1] = vel2.data_ptr()
    N.call("jrl_fluid_walk_f32", ctypes.byref(w), int(nsteps), like=f)
    N.launch_count += max(0, 3 * int(nsteps) - 1)
    try:
        torch.autograd.graph.increment_version(flags)
    except AttributeError:
        flags.add_(0)
    cur = int(w.current)
    return f, (rho2 if cur else rho), (vel2 if cur else vel), float(w.walks), phi, obs_vel


def new_sums(device):
    """Zeroed accumulator for FusedStep.step(sums=...)."""
    return torch.zeros((N.SUM_SLOTS, 4), dtype=torch.float64, device=device)


def total_sums(sums):
    """[mass, momentum_x, momentum_y, cells] as a device float64 [4] tensor (no host sync)."""
    return sums.sum(dim=0)
