"""jrl_step_noalloc_f32 -- the allocation-free variant of jrl_step_f32 (include/jrl_lbm.h; no counterpart in the reference)
-- against the existing entry point, through the C ABI (ctypes), on two
calls in a row.

Collected LAST on purpose (file name): a failure here must not hide the parity
tests.  Runs alone in a fresh process or after the others: it builds its own inputs and loads the library itself.

Both entry points run the same pipelined kernel (W >= 512, mixed flags, so the tile mask AND the boundary-group work
list are exercised); the existing one gets torch-allocated scratch, the new one uses the library's static device
array.  Tolerance: the suite's 1e-5 relative L2 (tests/util.py); the results are expected to be bit-identical, which
is reported, not required."""
import ctypes

import numpy as np
import pytest
import torch

from util import REL_L2_TOL, rel_l2


def _problem(dev, H=96, W=1024, seed=3):
    g = torch.Generator().manual_seed(seed)
    flags = torch.ones((1, 1, H, W), dtype=torch.uint8)
    r = torch.rand((1, 1, H, W), generator=g)
    flags[r < 0.02] = 2                      # scattered obstacles: boundary groups for the work list
    flags[:, :, 40:56, 300:340] = 2          # a block
    flags[:, :, 0, :] = 2
    flags[:, :, -1, :] = 2
    f = torch.rand((1, 9, H, W), generator=g) * 0.1 + 0.05
    rho = f.sum(dim=1, keepdim=True)
    vel = (torch.rand((1, 2, H, W), generator=g) - 0.5) * 0.05
    return flags.to(dev).contiguous(), f.to(dev).contiguous(), rho.to(dev).contiguous(), vel.to(dev).contiguous()


def _desc(N, flags, f_in, f_out, rho, vel, rho_out, vel_out):
    B, _, H, W = f_in.shape
    d = N.StepDesc()
    d.B, d.H, d.W = B, H, W
    d.axisym, d.op, d.kbc_type, d.is_convection = 0, N.OP_BGK, 0, 0
    d.dx, d.dt, d.tau = 1.0, 1.0, 0.8
    d.inflow_density = -1.0
    d.flags, d.f_in, d.f_out = flags.data_ptr(), f_in.data_ptr(), f_out.data_ptr()
    d.rho, d.vel, d.rho_out, d.vel_out = rho.data_ptr(), vel.data_ptr(), rho_out.data_ptr(), vel_out.data_ptr()
    return d


@pytest.mark.gpu
def test_noalloc_step_matches_step_twice_in_a_row():
    import jrl_native as N
    lib = N.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    flags, f0, rho0, vel0 = _problem(dev)
    B, _, H, W = f0.shape
    assert int(lib.jrl_step_noalloc_max_bytes()) == 4 << 20

    def run(noalloc):
        f, rho, vel = f0.clone(), rho0.clone(), vel0.clone()
        outs = []
        if not noalloc:  # the existing entry point with caller-allocated scratch
            mask = torch.empty((int(lib.jrl_tile_mask_bytes(B, H, W)),), dtype=torch.uint8, device=dev)
            wl = torch.zeros((int(lib.jrl_worklist_bytes(B, H, W)) // 4 + 1,), dtype=torch.int32, device=dev)
            assert lib.jrl_tile_mask_u8(ctypes.c_void_p(flags.data_ptr()), ctypes.c_void_p(mask.data_ptr()), B, H, W, stream) == 0
        for _ in range(2):  # two calls in a row: the second consumes the first one's output
            f2, rho2, vel2 = torch.empty_like(f), torch.empty_like(rho), torch.empty_like(vel)
            d = _desc(N, flags, f, f2, rho, vel, rho2, vel2)
            if noalloc:
                rc = lib.jrl_step_noalloc_f32(ctypes.byref(d), stream)
            else:
                d.tile_mask, d.worklist, d.worklist_bytes = mask.data_ptr(), wl.data_ptr(), wl.numel() * 4
                rc = lib.jrl_step_f32(ctypes.byref(d), stream)
            assert rc == 0, lib.jrl_last_error()
            torch.cuda.synchronize(dev)
            f, rho, vel = f2, rho2, vel2
            outs.append((f.cpu().numpy(), rho.cpu().numpy(), vel.cpu().numpy()))
        return outs

    want, got = run(False), run(True)
    exact = True
    for call, (w, g) in enumerate(zip(want, got)):
        for name, a, b in zip(("f", "rho", "vel"), w, g):
            assert np.isfinite(b).all(), f"call {call + 1}: {name} not finite"
            e = rel_l2(b, a)
            assert e <= REL_L2_TOL, f"call {call + 1}: {name} differs from jrl_step_f32 by {e:.3e} relative L2"
            exact = exact and np.array_equal(a, b)
    assert float(np.abs(want[1][0] - want[0][0]).max()) > 0, "the two calls must not be the same state"
    print("jrl_step_noalloc_f32 == jrl_step_f32 on two consecutive steps; bit-identical:", exact)


@pytest.mark.gpu
def test_noalloc_step_rejects_what_it_cannot_hold():
    """Sizes beyond the static scratch and caller-supplied scratch are refused before anything is launched."""
    import jrl_native as N
    lib = N.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    flags, f0, rho0, vel0 = _problem(dev, H=16, W=512)
    f2, rho2, vel2 = torch.empty_like(f0), torch.empty_like(rho0), torch.empty_like(vel0)
    d = _desc(N, flags, f0, f2, rho0, vel0, rho2, vel2)
    d.H, d.W = 8192, 8192                       # 8 MiB of work list: beyond the 4 MiB array (rejected, nothing launched)
    assert lib.jrl_step_noalloc_f32(ctypes.byref(d), stream) == -1
    assert b"too large" in lib.jrl_last_error()
    d = _desc(N, flags, f0, f2, rho0, vel0, rho2, vel2)
    d.tile_mask = flags.data_ptr()
    assert lib.jrl_step_noalloc_f32(ctypes.byref(d), stream) == -1
    assert b"must be NULL" in lib.jrl_last_error()
    d = _desc(N, flags, f0, f2, rho0, vel0, rho2, vel2)
    assert lib.jrl_step_noalloc_f32(ctypes.byref(d), stream) == 0, lib.jrl_last_error()
    torch.cuda.synchronize(dev)
