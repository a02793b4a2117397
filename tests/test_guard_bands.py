"""Own bounds / race evidence (compute-sanitizer is closed on this GPU pool; SURVEY section 5, VERDICT r1 missing #7).

  * guard bands: every output buffer of a step (f_out, rho_out, vel_out) is carved out of a larger allocation filled
    with NaN sentinels; after the launches no sentinel may have been touched -- an out-of-bounds store of the pipelined
    kernel (TMA ring, partial last tile), of the work-list pass or of the register kernel would show;
  * determinism: the same step from the same inputs, ten times, bit-identical every time -- a race on the mbarrier
    ring, on the work list or on the in-place INFLOW handling (the reference's own kernel races at OUTFLOW cells,
    propagation.cu:195-255) would show as run-to-run differences.
Mixed flags with every CellType, phi_obs, inflow parameters; the pipelined kernel (W = 1040: partial last tile), the
register kernels (W = 1001, one cell per thread; W = 258, two), BGK / MRT / KBC / concentration forms."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S

PAD = 4096  # elements (16 KiB): keeps the carved tensors 16-byte aligned


def _guarded(shape, dev):
    n = int(np.prod(shape))
    buf = torch.full((n + 2 * PAD,), float("nan"), dtype=torch.float32, device=dev)
    return buf, buf[PAD:PAD + n].view(*shape)


def _intact(buf, n):
    return bool(torch.isnan(buf[:PAD]).all()) and bool(torch.isnan(buf[PAD + n:]).all())


CASES = [("bgk", None, False, (24, 1040)), ("mrt", None, False, (24, 1040)), ("kbc", 133, False, (20, 1040)),
         ("bgk", None, True, (24, 1040)), ("bgk", None, False, (17, 1001)), ("kbc", 138, False, (18, 258))]


@pytest.mark.gpu
@pytest.mark.parametrize("op,kbc,conv,res", CASES)
def test_guard_bands_and_determinism(cuda_api, op, kbc, conv, res):
    import jrl_step
    dev = torch.device("cuda:0")
    H, W = res
    st = S.random_state(res, 9, dev)
    P = cuda_api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.8, contact_angle=None,
                                      axisymmetric_type=1 if conv else 0)
    fs = jrl_step.FusedStep(P, operator=op, kbc_type=kbc, is_convection=conv, tau_D=0.9 if conv else None)
    mesh = cuda_api.create_2d_meshgrid_tensor(size=(1, 1, H, W), device=dev, dtype=torch.float32) + 0.5
    inflow_vel = torch.tensor([0.01, 0.02], device=dev)
    first = None
    for rep in range(10):
        bf, f_out = _guarded((1, 9, H, W), dev)
        br, rho_out = _guarded((1, 1, H, W), dev)
        bv, vel_out = _guarded((1, 2, H, W), dev)
        kw = dict(f_out=f_out, rho_out=rho_out, phi_obs=st["phi"], inflow_density=1.0)
        if conv:
            out = fs.step(st["f"], st["flags"], st["rho"], st["vel"], mesh_grid=mesh, clamp01=True, **kw)
            got = [out[0], out[1]]
        else:
            out = fs.step(st["f"], st["flags"], st["rho"], st["vel"], vel_out=vel_out, inflow_vel=inflow_vel, force=st["force"], **kw)
            got = [out[0], out[1], out[2]]
        torch.cuda.synchronize()
        assert _intact(bf, f_out.numel()) and _intact(br, rho_out.numel()), "a step wrote outside its output buffers"
        assert conv or _intact(bv, vel_out.numel()), "a step wrote outside vel_out"
        got = [g.clone() for g in got]
        if first is None:
            first = got
            assert all(torch.isfinite(g[:, :, 1:-1, 1:-1]).any() for g in got)
        else:
            for a, b in zip(first, got):
                assert torch.equal(a.view(torch.int32), b.view(torch.int32)), f"run {rep} differs from run 0: a race"
