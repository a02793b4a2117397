"""Test infrastructure (oracle/): run the REFERENCE on the CPU.  Not product code.

Two layers:

1. ``RefKernels`` -- ctypes binding of ``oracle/_ref/libjrl_ref.so`` (the reference's own
   CUDA kernel bodies compiled for the host by ``oracle/build_ref.sh``).  Works on numpy
   arrays; travels to the GPU box as a prebuilt .so.

2. ``import_reference()`` -- only usable where ``/root/reference`` exists (this container).
   It replaces ``torch.utils.cpp_extension.load`` by a factory that returns stand-in
   extension modules backed by ``RefKernels`` (same function names and argument order as
   the pybind modules: propagation.cpp:30-33, collision_2d.cpp:72-76, macro_compute.cpp:62-65,
   solid_fluid_coupling.cpp:35-38), stubs ``imageio``/``h5py`` when missing, and imports the
   reference's ``src.LBM`` package unmodified.  The reference's Python solver classes then
   run end to end on CPU tensors; ``oracle/gen_golden.py`` uses that to write
   ``tests/golden/*.npz``.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / ``--impl reference`` legs
may import this module.
"""
from __future__ import annotations

import ctypes
import os
import sys
import types

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libjrl_ref.so")


def have_ref_lib() -> bool:
    return os.path.exists(REF_LIB)


def _ptr(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.c_void_p)


class RefKernels:
    """numpy front-end to the reference kernel bodies compiled for the CPU."""

    def __init__(self, path: str = REF_LIB):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} missing: run oracle/build_ref.sh where /root/reference exists"
            )
        self.lib = ctypes.CDLL(path)
        L = self.lib
        vp, i32, f64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_double
        L.jrl_ref_prop.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp, f64, i32, i32, i32, i32, i32, i32]
        L.jrl_ref_grad.argtypes = [i32, vp, vp, vp, f64, i32, i32, i32, i32]
        L.jrl_ref_div.argtypes = [i32, vp, vp, vp, f64, i32, i32, i32, i32]
        L.jrl_ref_macro_C.argtypes = [i32, f64, f64, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32]
        L.jrl_ref_macro.argtypes = [i32, f64, f64, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32]
        L.jrl_ref_wall.argtypes = [i32, f64, f64, f64, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32]
        L.jrl_ref_num_threads.restype = i32
        L.jrl_ref_set_num_threads.argtypes = [i32]

    @property
    def num_threads(self) -> int:
        return int(self.lib.jrl_ref_num_threads())

    def set_num_threads(self, n: int) -> None:
        self.lib.jrl_ref_set_num_threads(int(n))

    @staticmethod
    def _isd(a) -> int:
        if a.dtype == np.float64:
            return 1
        assert a.dtype == np.float32, a.dtype
        return 0

    # K1, propagation.cu:22-488.  rho / vel are mutated in place like the original.
    def prop(self, flags, f, phi_obs=None, rho=None, vel=None, inflow_vel=None,
             inflow_density=-1.0, is_convection=False, axisym=0, outflow_last=True, f_new=None):
        B, Q, H, W = f.shape
        assert Q == 9 and flags.dtype == np.uint8
        if f_new is None:
            f_new = f.copy()  # LBM_propagation_2d.py:49  (f_new = f + 0.0)
        self.lib.jrl_ref_prop(self._isd(f), _ptr(flags), _ptr(f), _ptr(phi_obs), _ptr(f_new), _ptr(rho),
                              _ptr(vel), _ptr(inflow_vel), float(inflow_density), int(is_convection),
                              int(axisym), B, H, W, int(outflow_last))
        return f_new

    # K2, collision_2d.cu:22-148
    def grad(self, rho, flags, dx=1.0, axisym=0):
        B, _, H, W = rho.shape
        out = np.zeros((B, 2, H, W), rho.dtype)
        self.lib.jrl_ref_grad(self._isd(rho), _ptr(rho), _ptr(flags), _ptr(out), float(dx), int(axisym), B, H, W)
        return out

    # K3, collision_2d.cu:150-220
    def div(self, vel, flags, dx=1.0, axisym=0):
        B, _, H, W = vel.shape
        out = np.zeros((B, 1, H, W), vel.dtype)
        self.lib.jrl_ref_div(self._isd(vel), _ptr(vel), _ptr(flags), _ptr(out), float(dx), int(axisym), B, H, W)
        return out

    # K5, macro_compute.cu:22-78
    def macro_C(self, dx, dt, h, C, flags, vel, mesh_grid, axisym=0):
        B, _, H, W = h.shape
        out = C.copy()
        self.lib.jrl_ref_macro_C(self._isd(h), float(dx), float(dt), _ptr(h), _ptr(C), _ptr(flags), _ptr(vel),
                                 _ptr(mesh_grid), _ptr(out), int(axisym), B, H, W)
        return out

    # K6, macro_compute.cu:80-157
    def macro(self, dx, dt, f, rho, vel, flags, density=None, axisym=0):
        B, _, H, W = f.shape
        rho_new, vel_new = rho.copy(), vel.copy()
        den_new = None if density is None else density.copy()
        self.lib.jrl_ref_macro(self._isd(f), float(dx), float(dt), _ptr(f), _ptr(rho), _ptr(vel), _ptr(flags),
                               _ptr(rho_new), _ptr(vel_new), _ptr(den_new), int(axisym), B, H, W)
        return (rho_new, vel_new) if den_new is None else (rho_new, vel_new, den_new)

    # K7, solid_fluid_coupling.cu:25-252
    def wall(self, dt, dx, tau, rho, vel, flags, f, phi_obs, obs_vel=None, is_convection=False, axisym=0):
        B, _, H, W = f.shape
        f_new = f.copy()
        self.lib.jrl_ref_wall(self._isd(f), float(dt), float(dx), float(tau), _ptr(rho), _ptr(vel), _ptr(flags),
                              _ptr(f), _ptr(f_new), _ptr(phi_obs), _ptr(obs_vel), int(is_convection),
                              int(axisym), B, H, W)
        return f_new


# ----------------------------------------------------------------------------------------
# Importing the reference's Python package on the CPU (this container only).
# ----------------------------------------------------------------------------------------

def _t2n(t):
    """torch CPU tensor -> numpy view sharing memory, or None for an empty tensor
    (the reference passes `torch.Tensor([])` for absent optionals, propagation.cu:513-517)."""
    if t is None or t.numel() == 0:
        return None
    assert t.device.type == "cpu" and t.is_contiguous()
    return t.detach().numpy()


def _make_extension(name: str, ker: RefKernels, outflow_last: bool):
    m = types.ModuleType(name)
    if name == "propagation":
        def propagation_2d_forward(flags, f, phi_obs, f_new, rho, vel, inflow_vel, inflow_density,
                                   is_convection, axisymmetric_type):
            ker.prop(_t2n(flags), _t2n(f), _t2n(phi_obs), _t2n(rho), _t2n(vel), _t2n(inflow_vel),
                     inflow_density, is_convection, axisymmetric_type, outflow_last, f_new=_t2n(f_new))
        m.propagation_2d_forward = propagation_2d_forward
    elif name == "collision_2d":
        def get_grad_2d_forward(rho, flags, grad_rho, dx, axisymmetric_type):
            _t2n(grad_rho)[...] = ker.grad(_t2n(rho.contiguous()), _t2n(flags), dx, axisymmetric_type)

        def get_div_2d_forward(vel, flags, div_vel, dx, axisymmetric_type):
            _t2n(div_vel)[...] = ker.div(_t2n(vel.contiguous()), _t2n(flags), dx, axisymmetric_type)
        m.get_grad_2d_forward = get_grad_2d_forward
        m.get_div_2d_forward = get_div_2d_forward
        m.collision_2d_forward = get_div_2d_forward  # sic: collision_2d.cpp:75 binds it so
    elif name == "macro_compute":
        def macro_compute_C_2d_forward(dx, dt, h, C, flags, vel, mesh_grid, C_new, axisymmetric_type):
            _t2n(C_new)[...] = ker.macro_C(dx, dt, _t2n(h), _t2n(C), _t2n(flags), _t2n(vel), _t2n(mesh_grid),
                                           axisymmetric_type)

        def macro_compute_2d_forward(dx, dt, f, rho, vel, flags, rho_new, vel_new, g, pressure, density,
                                     density_new, axisymmetric_type):
            res = ker.macro(dx, dt, _t2n(f), _t2n(rho), _t2n(vel), _t2n(flags), _t2n(density), axisymmetric_type)
            _t2n(rho_new)[...] = res[0]
            _t2n(vel_new)[...] = res[1]
            if len(res) == 3:
                _t2n(density_new)[...] = res[2]
        m.macro_compute_C_2d_forward = macro_compute_C_2d_forward
        m.macro_compute_2d_forward = macro_compute_2d_forward
    elif name == "solid_fluid_coupling":
        def solve_obstacle_2d_forward(dt, dx, tau, rho, vel, flags, f, f_new, phi_obs, obs_vel,
                                      is_convection, axisymmetric_type):
            _t2n(f_new)[...] = ker.wall(dt, dx, tau, _t2n(rho), _t2n(vel), _t2n(flags), _t2n(f), _t2n(phi_obs),
                                        _t2n(obs_vel), is_convection, axisymmetric_type)
        m.solve_obstacle_2d_forward = solve_obstacle_2d_forward
    else:
        raise KeyError(name)
    return m


def import_reference(ref_root: str = "/root/reference", outflow_last: bool = True):
    """Import the reference's `src.LBM` on the CPU. Returns the `src.LBM` module tree root
    (`sys.modules['src']`). Must run in a process that has not imported this repo's own
    `src.LBM` drop-in (same package name by design)."""
    import torch
    import torch.utils.cpp_extension as cpp_ext

    if not os.path.isdir(os.path.join(ref_root, "src", "LBM")):
        raise FileNotFoundError(f"no reference tree at {ref_root}")
    if "src" in sys.modules and not getattr(sys.modules["src"], "__jrl_reference__", False):
        raise RuntimeError("a different `src` package is already imported in this process")
    ker = RefKernels()
    cpp_ext.load = lambda name, **kw: _make_extension(name, ker, outflow_last)  # noqa: E731
    standins = os.path.join(os.path.dirname(_HERE), "tests", "standins")
    for missing in ("imageio", "h5py"):
        try:
            __import__(missing)
        except ImportError:
            if os.path.isdir(standins) and standins not in sys.path:
                sys.path.append(standins)  # tests/standins: the few calls the demos make after their loops
            try:
                __import__(missing)
            except ImportError:
                sys.modules[missing] = types.ModuleType(missing)
    sys.path.insert(0, ref_root)
    import src.LBM.simulation  # noqa: F401  (pulls in all four solver packages)
    import src
    src.__jrl_reference__ = True
    torch.set_num_threads(max(1, torch.get_num_threads()))
    return src
