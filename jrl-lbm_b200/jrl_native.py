"""ctypes binding of libjrl_lbm.so (include/jrl_lbm.h) -- the only way this package computes.

There is NO CPU fallback: `load()` opens the nvcc-built library that sits next to this file
and raises if it is missing, and every wrapper rejects tensors that are not CUDA / contiguous
/ fp32 (uint8 for flags).  PyTorch is used for device memory and streams only; pointers are
`tensor.data_ptr()` and the stream is `torch.cuda.current_stream()`.

(The CPU test-suite exercises the same kernel SOURCES compiled for the host by
tests/host_emu/build.sh; tests/host_emu/install.py monkeypatches this module from the outside to do
so.  Nothing in this package knows about that library.)
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libjrl_lbm.so")

_lib = None

AXI_NOT, AXI_X, AXI_Y = 0, 1, 2
TIP_BOX, TIP_CONE45 = 0, 1
OP_BGK, OP_MRT, OP_KBC, OP_NONE = 0, 1, 2, 3
SUM_SLOTS = 64


class CollideDesc(Structure):  # struct jrl_collide_desc
    _fields_ = [("B", c_int), ("H", c_int), ("W", c_int), ("axisym", c_int), ("op", c_int), ("kbc_type", c_int),
                ("is_convection", c_int), ("dx", c_double), ("dt", c_double), ("tau", c_double),
                ("mrt", c_void_p), ("f", c_void_p), ("rho", c_void_p), ("vel", c_void_p), ("flags", c_void_p),
                ("force", c_void_p), ("mesh_grid", c_void_p), ("f_out", c_void_p), ("slab_row0", c_int), ("slab_rows", c_int)]


class StepDesc(Structure):  # struct jrl_step_desc
    _fields_ = [("B", c_int), ("H", c_int), ("W", c_int), ("axisym", c_int), ("op", c_int), ("kbc_type", c_int),
                ("is_convection", c_int), ("use_contact", c_int), ("contact_tan", c_float), ("clamp01", c_int),
                ("dx", c_double), ("dt", c_double), ("tau", c_double), ("mrt", c_void_p), ("flags", c_void_p),
                ("f_in", c_void_p), ("f_out", c_void_p), ("phi_obs", c_void_p), ("rho", c_void_p),
                ("vel", c_void_p), ("rho_out", c_void_p), ("vel_out", c_void_p), ("force", c_void_p),
                ("mesh_grid", c_void_p), ("inflow_vel", c_void_p), ("inflow_density", c_double),
                ("use_wall", c_int), ("obs_vel", c_void_p), ("slab_row0", c_int), ("slab_rows", c_int),
                ("row_begin", c_int), ("row_end", c_int), ("force_vec", c_int), ("sums", c_void_p), ("wall_phi", c_void_p),
                ("worklist", c_void_p), ("worklist_bytes", ctypes.c_longlong), ("tile_mask", c_void_p)]


class TipDesc(Structure):  # struct jrl_tip_desc
    _fields_ = [("H", c_int), ("W", c_int), ("shape", c_int), ("is_negative", c_int), ("walks_int", c_int), ("Rg_radius", c_int),
                ("radius_obs_int", c_int), ("box_cx", c_float), ("box_cy", c_float), ("box_radius", c_float),
                ("vel_obs_conv", c_float), ("w0", c_float), ("w1", c_float), ("mesh_grid", c_void_p), ("flags", c_void_p),
                ("C", c_void_p), ("vel0", c_void_p), ("vel1", c_void_p), ("phi_obs", c_void_p), ("C_vel", c_void_p)]


class TipFluidDesc(Structure):  # struct jrl_tip_fluid_desc
    _fields_ = [("H", c_int), ("W", c_int), ("shape", c_int), ("walks_int", c_int), ("Rg_radius", c_int), ("box_cx", c_float), ("box_cy", c_float),
                ("box_radius", c_float), ("vel_obs", c_float), ("mesh_grid", c_void_p), ("flags", c_void_p), ("vel", c_void_p),
                ("phi_obs", c_void_p), ("obs_vel", c_void_p)]


class FluidWalkDesc(Structure):  # struct jrl_fluid_walk_desc
    _fields_ = [("collide", CollideDesc), ("tip", TipFluidDesc), ("walks", c_double), ("vel_obs", c_double),
                ("box_radius", c_double), ("f", c_void_p * 3), ("rho", c_void_p * 2), ("vel", c_void_p * 2), ("current", c_int)]


class WalkDesc(Structure):  # struct jrl_walk_desc
    _fields_ = [("step", StepDesc), ("tip", TipDesc), ("walks", c_double), ("vel_obs_conv", c_double), ("dx", c_double),
                ("dt", c_double), ("box_radius", c_double), ("inflow_height", c_double), ("x_interval", c_double),
                ("conv_timer", c_double), ("dt_conv_real", c_double), ("vel_obs_real", c_double), ("radius_obs_real", c_double),
                ("radius_obs", c_double), ("steps_done", c_int),
                ("vel_snaps", c_void_p), ("n_snaps", c_int), ("records", c_void_p)]


_vp, _i, _d, _f = c_void_p, c_int, c_double, c_float
# name -> argtypes; every symbol include/jrl_lbm.h declares (tests/test_abi.py checks the list against the header)
PROTOTYPES = {
    "jrl_version": [],
    "jrl_last_error": [],
    "jrl_stream_f32": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _d, _i, _i, _i, _i, _vp],
    "jrl_inflow_apply_f32": [_vp, _vp, _vp, _vp, _d, _i, _i, _i, _vp],
    "jrl_rebounce_f32": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "jrl_moments_f32": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _d, _i, _i, _i, _vp],
    "jrl_moments_kernel_f32": [_vp, _vp, _vp, _vp, _vp, _vp, _d, _i, _i, _i, _vp],
    "jrl_moments_C_f32": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "jrl_moments_C_kernel_f32": [_d, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "jrl_feq_f32": [_vp, _vp, _vp, _vp, _d, _d, _i, _i, _i, _vp],
    "jrl_grad_f32": [_vp, _vp, _vp, _d, _i, _i, _i, _i, _vp],
    "jrl_div_f32": [_vp, _vp, _vp, _d, _i, _i, _i, _i, _vp],
    "jrl_collide_f32": [POINTER(CollideDesc), _vp],
    "jrl_wall_f32": [_d, _d, _d, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "jrl_strip_flux_f32": [_vp, _vp, _i, _i, _vp, _i, _i, _vp],
    "jrl_momentum_exchange_f32": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "jrl_step_f32": [POINTER(StepDesc), _vp],
    "jrl_step_noalloc_f32": [POINTER(StepDesc), _vp],
    "jrl_step_noalloc_max_bytes": [],
    "jrl_run_f32": [POINTER(StepDesc), _i, _i, ctypes.c_longlong, _vp],
    "jrl_tile_mask_bytes": [_i, _i, _i],
    "jrl_worklist_bytes": [_i, _i, _i],
    "jrl_tile_mask_u8": [_vp, _vp, _i, _i, _i, _vp],
    "jrl_tip_prepare_conc_f32": [POINTER(TipDesc), _vp],
    "jrl_tip_prepare_fluid_f32": [POINTER(TipFluidDesc), _vp],
    "jrl_fluid_walk_f32": [POINTER(FluidWalkDesc), _i, _vp],
    "jrl_conc_walk_f32": [POINTER(WalkDesc), _i, _vp],
    "jrl_halo_pack_f32": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "jrl_halo_unpack_f32": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
}


def _bind(lib):
    for name, argtypes in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.argtypes = argtypes
        fn.restype = c_char_p if name == "jrl_last_error" else (ctypes.c_longlong if name in ("jrl_worklist_bytes", "jrl_step_noalloc_max_bytes") else c_int)
    return lib


def load():
    """Open libjrl_lbm.so (built by `make -C jrl-lbm_b200` / __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build the sm_100a library first (make -C jrl-lbm_b200). "
                "This package has no CPU or PyTorch fallback.")
        _lib = _bind(ctypes.CDLL(LIB_PATH))
    return _lib


def _require_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name}: tensor is on {t.device}; JRL-LBM-B200 has no CPU path (CUDA tensors only)")


def stream_ptr(like: torch.Tensor):
    if like.is_cuda:
        return c_void_p(torch.cuda.current_stream(like.device).cuda_stream)
    return None


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().jrl_last_error()
        raise RuntimeError(f"libjrl_lbm {what}: {msg.decode() if msg else rc}")


def tensor_ptr(t, name, dtype=torch.float32, shape=None, like=None, optional=False):
    """Validated raw device pointer of `t` (the reference checks nothing and
    reinterprets memory on a wrong dtype; the drop-in raises instead)."""
    if t is None or (isinstance(t, torch.Tensor) and t.numel() == 0 and optional):
        if optional:
            return None
        raise RuntimeError(f"{name}: tensor required")
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f"{name}: expected a torch.Tensor, got {type(t).__name__}")
    if t.dtype != dtype:
        raise RuntimeError(f"{name}: dtype {t.dtype} not supported, expected {dtype} (fp32 path only)")
    _require_cuda(t, name)
    if like is not None and t.device != like.device:
        raise RuntimeError(f"{name}: device {t.device} differs from {like.device}")
    if not t.is_contiguous():
        raise RuntimeError(f"{name}: tensor must be contiguous [B,C,H,W]")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise RuntimeError(f"{name}: shape {tuple(t.shape)}, expected {tuple(shape)}")
    return c_void_p(t.data_ptr())


class device_of:
    """Context manager: make `t`'s device current so the launch lands on it (multi-GPU processes)."""

    def __init__(self, t):
        self.ctx = torch.cuda.device(t.device) if t.is_cuda else None

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


# count of library launches issued through this module (bench.py's gpu_launches claim)
launch_count = 0


def call(name, *args, like):
    """Invoke one library entry point on `like`'s device and current stream."""
    global launch_count
    lib = load()
    with device_of(like):
        rc = getattr(lib, name)(*args, stream_ptr(like))
    launch_count += 1
    check(rc, name)
