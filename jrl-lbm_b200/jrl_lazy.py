"""The fused lattice update behind the reference's FOUR calls (SURVEY 7.3 hard part 1).

The reference's demos drive one time step as

    f          = prop.propagation(f, flags, ...)                 # K1
    f          = coupling.solve_boundary(..., f=f, ...)          # K7 (cylinder / JRL fluid demos only)
    rho, vel   = macro.macro_compute(f=f, rho=rho, vel=vel, ...) # or C = macro.macro_compute_C(h=f, ...)
    f          = collision.collision(f=f, rho=rho, vel=vel, ...)

(demo_2d_LBM.py:101-119, demo_2d_LBM_cylinder.py:138-172, demo_2d_LBM_JRL_C_cylinder.py:313-340).  Run one library
launch per call that is four passes over the populations (72 + 72 + 48 + 84 B per cell); the fused kernel
(`jrl_step_f32`) does the same arithmetic in ONE pass at 72 B per cell.  This module lets the unchanged demo loop reach
it: the first three calls only RECORD what they were asked (a `PendingStep`) and hand back `LazyField`s -- tensor
subclasses without storage that know their shape / dtype / device -- and `collision()` launches the fused step when it
is given exactly the chain it recognises, filling in `rho` / `vel` (`C`) for the caller.  Results are bit-identical to
the four launches (tests/test_lazy_fusion.py); `fused == chain` is a property of the kernels.

Nothing is ever guessed: the moment ANY torch operation touches a LazyField (indexing, arithmetic, `.cpu()`, printing,
in-place writes: `__torch_function__` sees them all) the recorded calls are executed eagerly, in the reference's order,
one launch each, and the field behaves as the ordinary tensor from then on.  Streaming's in-place side effects on
`rho` / `vel` at INFLOW cells (propagation.cu:300-362) are kept: the tensors `propagation()` returns for them are
LazyFields too, and touching one runs the streaming launch that writes them.  A recorded input that is mutated before
the step was launched (its torch version counter moved) cannot be replayed; that raises instead of computing something
else.  `JRL_LBM_FUSE=0` in the environment turns the whole mechanism off (every call launches immediately).
"""
from __future__ import annotations

import os
import sys
import weakref

import torch
from torch.utils._pytree import tree_map_only

ENABLED = os.environ.get("JRL_LBM_FUSE", "1") != "0"

T = torch.Tensor
_META = {
    T.shape.__get__, T.dtype.__get__, T.device.__get__, T.ndim.__get__, T.is_cuda.__get__, T.layout.__get__,
    T.requires_grad.__get__, T.grad_fn.__get__, T.is_leaf.__get__, T.is_sparse.__get__, T.is_quantized.__get__,
    T.dim, T.size, T.numel, T.nelement, T.is_contiguous, T.stride, T.is_floating_point, T.is_complex, T.element_size,
    T.get_device, T.__len__, T.ndimension, T.storage_offset,
}

# statistics for the tests and tools: how many steps were fused / replayed eagerly, and why
stats = {"fused": 0, "eager": 0, "reasons": {}}


def _why(reason):
    stats["eager"] += 1
    stats["reasons"][reason] = stats["reasons"].get(reason, 0) + 1


class LazyField(torch.Tensor):
    """A field of a recorded-but-not-yet-launched lattice update.  `real()` returns the tensor, launching if needed."""

    @staticmethod
    def __new__(cls, shape, dtype, device, pending, role):
        r = torch.Tensor._make_wrapper_subclass(cls, tuple(shape), dtype=dtype, device=device, requires_grad=False)
        r._p, r._role, r._real = pending, role, None
        return r

    def real(self):
        if self._real is None:
            self._p.materialise(self._role)
        return self._real

    def _set(self, t):
        self._real, self._p = t, None  # a resolved field no longer keeps the recorded inputs alive

    @property
    def pending(self):
        """The PendingStep this field still depends on, or None once it is an ordinary tensor."""
        return self._p if self._real is None else None

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func in _META:
            with torch._C.DisableTorchFunctionSubclass():
                return func(*args, **kwargs)
        if func is torch.clamp and len(args) == 3 and not kwargs and isinstance(args[0], LazyField):
            p = args[0].pending  # `C = torch.clamp(C, 0, 1)` between macro_compute_C and collision
            if p is not None and args[0]._role == "C" and args[1] == 0 and args[2] == 1 and not p.clamp:
                return p.add_clamp()
        args, kwargs = tree_map_only(LazyField, lambda t: t.real(), (args, kwargs))
        return func(*args, **kwargs)

    @classmethod
    def __torch_dispatch__(cls, func, types, args=(), kwargs=None):
        # anything that reaches the dispatcher without passing __torch_function__ (C++ callers): same rule
        args, kwargs = tree_map_only(LazyField, lambda t: t.real(), (args, kwargs or {}))
        return func(*args, **kwargs)

    def __repr__(self):
        return repr(self.real())


def resolve(x):
    """x as an ordinary tensor (launching what it is waiting for); non-LazyFields pass through."""
    return x.real() if isinstance(x, LazyField) else x


def _ver(t):
    return None if t is None else (t, t._version)


class PendingStep:
    """One time step's recorded calls.  `ops` are the eager single-launch implementations (jrl_solver)."""

    def __init__(self, prop, f_in, flags, phi_obs, rho_in, vel_in, inflow_vel, inflow_density):
        self.prop = prop
        self.f_in, self.flags, self.phi_obs = f_in, flags, phi_obs
        self.rho_in, self.vel_in = rho_in, vel_in
        self.inflow_vel, self.inflow_density = inflow_vel, inflow_density
        self.wall = None       # dict(obj, dt, dx, tau, rho, vel, phi_obs, obs_vel)
        self.moments = None    # dict(obj, kind, dx, dt, rho, vel, mesh_grid)
        self.clamp = False
        self.real = {}         # role -> tensor, for the stages that have been executed
        self.handles = {}      # role -> weakref of the LazyField handed out
        self.launched = False  # the fused step ran
        self._watch = [_ver(t) for t in (f_in, flags, phi_obs, rho_in, vel_in, inflow_vel) if t is not None]

    # -- handles -------------------------------------------------------------------------------------------
    def handle(self, role, shape, like):
        h = LazyField(shape, like.dtype, like.device, self, role)
        self.handles[role] = weakref.ref(h)
        return h

    def _publish(self, role, t):
        self.real[role] = t
        ref = self.handles.get(role)
        h = ref() if ref is not None else None
        if h is not None and h._real is None:
            h._set(t)

    def owns(self, x, *roles):
        """x is an unresolved field of this step with one of `roles`."""
        return isinstance(x, LazyField) and x._real is None and x._p is self and x._role in roles

    def is_input(self, x, original, inplace_role):
        """x denotes the recorded input `original` (the tensor itself, or the in-place handle propagation returned)."""
        if x is original:
            return True
        if isinstance(x, LazyField):
            return (x._real is original) or (x._real is None and x._p is self and x._role == inplace_role)
        return False

    def watch(self, *tensors):
        self._watch += [_ver(t) for t in tensors if isinstance(t, torch.Tensor) and not isinstance(t, LazyField)]

    def check_unmodified(self):
        for t, v in self._watch:
            if t._version != v:
                names = [n for n in ("f_in", "flags", "phi_obs", "rho_in", "vel_in", "inflow_vel") if getattr(self, n) is t]
                raise RuntimeError(
                    f"[{', '.join(names) or 'a later argument'}: version {v} -> {t._version}] "
                    "JRL-LBM-B200: a tensor passed to propagation() / solve_boundary() / macro_compute() was modified "
                    "in place before the step was launched; the deferred step cannot be replayed.  Set JRL_LBM_FUSE=0 "
                    "to launch every call immediately.")

    # -- recording -----------------------------------------------------------------------------------------
    def add_wall(self, obj, dt, dx, tau, rho, vel, phi_obs, obs_vel):
        # `rho` is accepted and ignored by the reference's kernel too; not kept (it may be propagation()'s handle)
        self.wall = dict(obj=obj, dt=dt, dx=dx, tau=tau, rho=None, vel=vel, phi_obs=phi_obs, obs_vel=obs_vel)
        self.watch(vel, phi_obs, obs_vel)
        return self.handle("f_wall", self.f_in.shape, self.f_in)

    def add_moments(self, obj, kind, dx, dt, rho, vel, mesh_grid=None):
        self.moments = dict(obj=obj, kind=kind, dx=dx, dt=dt, rho=rho, vel=vel, mesh_grid=mesh_grid)
        self.watch(rho, vel)
        B, _, H, W = self.f_in.shape
        if kind == "C":
            return self.handle("C", (B, 1, H, W), self.f_in)
        return [self.handle("rho", (B, 1, H, W), self.f_in), self.handle("vel", (B, 2, H, W), self.f_in)]

    def add_clamp(self):
        self.clamp = True
        B, _, H, W = self.f_in.shape
        return self.handle("C_clamp", (B, 1, H, W), self.f_in)

    def f_role(self):
        return "f_wall" if self.wall is not None else "f_stream"

    # -- eager replay (one launch per recorded call, the reference's order) ----------------------------------
    def materialise(self, role):
        if role in self.real:
            return self._publish(role, self.real[role])
        if role in ("f_stream", "rho_in", "vel_in"):
            if "f_stream" not in self.real:
                self.check_unmodified()
                _why("field touched before collision()" if not self.launched else "stale field touched after the fused step")
                f_new = self.prop._stream(self.f_in, self.flags, self.phi_obs, self.rho_in, self.vel_in, self.inflow_vel,
                                          self.inflow_density)
                self._publish("f_stream", f_new)
                if self.rho_in is not None:
                    self._publish("rho_in", self.rho_in)   # mutated in place by that launch
                if self.vel_in is not None:
                    self._publish("vel_in", self.vel_in)
            return
        if role == "f_wall":
            self.materialise("f_stream")
            w = self.wall
            self._publish("f_wall", w["obj"]._wall(w["dt"], w["dx"], w["tau"], resolve(w["vel"]), self.flags,
                                                   self.real["f_stream"], w["phi_obs"], w["obs_vel"]))
            return
        if role in ("rho", "vel", "C"):
            self.materialise(self.f_role())
            m, f = self.moments, self.real[self.f_role()]
            if m["kind"] == "C":
                self._publish("C", m["obj"]._moments_C(f, resolve(m["rho"]), self.flags))
            else:
                r, v = m["obj"]._moments(m["dx"], m["dt"], f, resolve(m["rho"]), resolve(m["vel"]), self.flags)
                self._publish("rho", r)
                self._publish("vel", v)
            return
        if role == "C_clamp":
            self.materialise("C")
            self._publish("C_clamp", torch.clamp(self.real["C"], 0, 1))
            return
        raise KeyError(role)

    def _visible_outside(self, role):
        """The recorded input `role` ('rho_in' / 'vel_in') is still reachable by the caller: its handle is alive and
        unresolved, or the tensor object has more references than this record's own."""
        ref = self.handles.get(role)
        h = ref() if ref is not None else None
        if h is not None and h._real is None:
            return True
        t = getattr(self, role)
        own = 1                                                          # the attribute
        own += sum(1 for w in self._watch if w[0] is t)                  # version watch list
        own += sum(1 for d in (self.moments, self.wall) if d for v in d.values() if v is t)
        return sys.getrefcount(t) - 2 > own                              # - getrefcount's argument and the local `t`

    # -- the fused launch ----------------------------------------------------------------------------------------
    def publish_fused(self, rho_out, vel_out):
        """After jrl_step_f32: the moments handles become the launch's outputs; the streamed / wall-corrected
        populations were never stored (their handles replay eagerly if anyone still looks at them)."""
        self.launched = True
        stats["fused"] += 1
        # jrl_step_f32 does not write the previous-step rho / vel.  If the caller can still see them -- through the
        # handles propagation() returned for them, or through the ORIGINAL tensor objects (more Python references than
        # this record itself holds) -- give them streaming's in-place writes now (one small launch).  The demos rebind
        # both names before collision(), so for them this never runs.  (A view of the original tensor held elsewhere
        # is not detected: set JRL_LBM_FUSE=0 for code that reads pre-step rho / vel through views after a step.)
        # streaming writes rho only with an inflow density (INFLOW cells) or at INFLOW_2 cells (rho = 1), and vel only
        # with an inflow velocity (propagation.cu:300, 320-321, 362)
        written = {"rho_in": self.inflow_density is not None or self.prop._has_inflow2(self.flags),
                   "vel_in": self.inflow_vel is not None}
        alive = [r for r in ("rho_in", "vel_in") if written[r] and getattr(self, r) is not None and self._visible_outside(r)]
        if alive:
            self.prop._inflow_apply(self.flags, self.rho_in if "rho_in" in alive else None,
                                    self.vel_in if "vel_in" in alive else None, self.inflow_vel, self.inflow_density)
        for r in alive:
            if r in self.handles:
                self._publish(r, self.rho_in if r == "rho_in" else self.vel_in)
        if self.moments["kind"] == "C":
            # with the clamp fused in, the unclamped sum was never stored: its handle replays eagerly if touched
            self._publish("C_clamp" if self.clamp else "C", rho_out)
        else:
            self._publish("rho", rho_out)
            self._publish("vel", vel_out)
