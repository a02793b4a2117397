"""GPU (-m gpu): the CUDA library, called through the drop-in `src.LBM` API and the C ABI, against
  (a) the golden fixtures written by the REFERENCE's own Python classes (oracle/gen_golden.py),
  (b) the numpy oracle on the same seeded inputs,
  (c) size-independent properties at bandwidth-bound sizes (fused == operator chain bit for bit,
      mass conservation on the periodic all-FLUID grid).
Tolerance: BASELINE.json north_star -- masks / integer outputs bit-exact, float fields within
1e-5 relative L2 (util.REL_L2_TOL)."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S
from fused_cases import (FUSED_CASES, RUN_CASES, assert_identical, check_run_equals_steps, check_device_tip, check_fluid_walk, check_run_sums_series,
                         check_sums, check_tile_mask, run_pair)
from golden_cases import CASES
from util import compare_outputs, load_golden, rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("name", list(CASES))
def test_case_matches_reference_golden(cuda_api, dev, name):
    fn, kw = CASES[name]
    got = getattr(S, fn)(cuda_api, dev, **kw)
    compare_outputs(got, load_golden(name), case=name)


@pytest.mark.parametrize("name", list(CASES))
def test_case_matches_oracle(cuda_api, oracle_api, dev, name):
    fn, kw = CASES[name]
    got = getattr(S, fn)(cuda_api, dev, **kw)
    want = getattr(S, fn)(oracle_api, torch.device("cpu"), **kw)
    compare_outputs(got, {k: v for k, v in want.items() if v is not None})


@pytest.mark.parametrize("name", list(FUSED_CASES))
def test_fused_equals_chain(cuda_api, dev, name):
    chain, fused = run_pair(cuda_api, dev, name)
    assert_identical(chain, fused)


@pytest.mark.parametrize("name", ["c1_shear_bgk", "c2_shear_mrt", "c3_cylinder", "c4_jrl_conc"])
def test_fused_matches_reference_golden(cuda_api, dev, name):
    fn, kw = CASES[name]
    got = getattr(S, fn)(cuda_api, dev, fused=True, **kw)
    skip = [k for k in load_golden(name) if k.startswith("snap_")]
    compare_outputs(got, load_golden(name), skip=skip, case=name)


def test_large_grid_fused_equals_chain_and_conserves_mass(cuda_api, dev):
    """2048 x 2048 periodic shear layer (larger than L2 with both buffers: 2 x 151 MB): the fused
    128-bit path equals the operator chain bit for bit, and total mass is conserved."""
    kw = dict(res=(2048, 2048), steps=6, operator="bgk")
    chain = S.shear_layer(cuda_api, dev, **kw)
    fused = S.shear_layer(cuda_api, dev, fused=True, **kw)
    for k in ("f", "rho", "vel"):
        assert np.array_equal(chain[k], fused[k]), k
    m0 = chain["f0"].astype(np.float64).sum()
    m1 = fused["f"].astype(np.float64).sum()
    assert abs(m1 - m0) / m0 < 1e-6


def test_vector_widths_agree(cuda_api, dev):
    """force_vec = 1 / 2 / 4 cells per thread give identical populations."""
    import jrl_step
    P = cuda_api.SimulationParameters(simulation_size=(1, 1, 96, 128), device=dev, tau=0.6, contact_angle=None)
    st = S.random_state((96, 128), 7, dev, obstacles=True, io_cells=True)
    fs = jrl_step.FusedStep(P, operator="bgk")
    iv = torch.tensor([0.01, 0.02], device=dev)
    outs = []
    for v in (1, 2, 4):
        f, r, u = fs.step(st["f"], st["flags"], st["rho"], st["vel"], phi_obs=st["phi"], force=st["force"],
                          inflow_vel=iv, inflow_density=1.0, force_vec=v)
        outs.append((f.cpu().numpy(), r.cpu().numpy(), u.cpu().numpy()))
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("op,conv,axisym", [("bgk", False, 0), ("mrt", False, 0), ("kbc", False, 0), ("bgk", True, 1),
                                            ("bgk", True, 0), ("bgk", True, 2)])
@pytest.mark.parametrize("W", [1024, 2064])
def test_pipelined_kernel_equals_register_kernel(cuda_api, dev, op, conv, axisym, W):
    """force_vec = 8 (TMA-pipelined kernel: 1024-cell tiles, partial last tile at W = 2064, halo wrap at
    the row ends, mixed FLUID / obstacle / inlet groups) equals force_vec = 4 and 1 bit for bit."""
    import jrl_step
    H = 37
    P = cuda_api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.6, contact_angle=None,
                                      axisymmetric_type=axisym)
    st = S.random_state((H, W), 9, dev, obstacles=True, io_cells=True)
    flags = st["flags"].clone()
    flags[:, :, 8:30, 100:W - 100] = 1          # a large all-FLUID region so that the staged (bulk) path runs too
    flags[:, :, 12:26, 0:24] = 1                # ... including groups on the x symmetry axis (column 0)
    flags[:, :, 9:20, 480:560] = 2              # a solid block (all-OBSTACLE groups: the pipelined kernel's data-movement
    flags[:, :, 0:3, 600:700] = 2               # fast path), across a tile boundary, and solid cells on the first / last
    flags[:, :, H - 2:H, 300:420] = 2           # rows and columns, where the mirrored neighbour rule applies (generic path)
    flags[:, :, 20:30, 0:12] = 2
    flags[:, :, 20:30, W - 10:W] = 2
    fs = jrl_step.FusedStep(P, operator=op, kbc_type=133 if op == "kbc" else None, is_convection=conv,
                            tau_D=0.8 if conv else None)
    mesh = cuda_api.create_2d_meshgrid_tensor((1, 1, H, W), device=dev) + 0.5
    kw = dict(phi_obs=st["phi"], inflow_density=0.5)
    if conv:
        kw.update(mesh_grid=mesh, clamp01=True)
    else:
        kw.update(force=st["force"], inflow_vel=torch.tensor([0.01, 0.02], device=dev))
    outs = []
    for v in (8, 4, 1):
        o = fs.step(st["f"], flags, st["rho"], st["vel"], force_vec=v, **kw)
        outs.append([t.cpu().numpy() for t in o])
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("name", list(RUN_CASES))
def test_run_equals_steps(cuda_api, oracle_api, dev, name):
    check_run_equals_steps(cuda_api, dev, name, oracle=oracle_api)


@pytest.mark.parametrize("is_negative,fused", [(True, True), (False, True), (True, False)])
def test_device_tip_motion(cuda_api, oracle_api, dev, is_negative, fused):
    check_device_tip(cuda_api, dev, is_negative=is_negative, fused=fused, oracle=oracle_api)


@pytest.mark.parametrize("gravity", [2e-5, 0.0])
def test_fluid_walk(cuda_api, oracle_api, dev, gravity):
    check_fluid_walk(cuda_api, dev, gravity, oracle=oracle_api)


def test_run_sums_series(cuda_api, dev):
    check_run_sums_series(cuda_api, dev)


@pytest.mark.parametrize("res", [(9, 1040), (8, 4096), (7, 528)])
def test_tile_mask(cuda_api, dev, res):
    check_tile_mask(cuda_api, dev, res=res)


def test_pipelined_kernel_with_and_without_tile_mask(cuda_api, dev):
    """The tile mask only lets all-FLUID tiles skip their flag rows: populations, moments and in-kernel sums are
    identical with it, without it, and after the flags change under a cached mask (version counter)."""
    import jrl_step
    H, W = 24, 4096
    P = cuda_api.SimulationParameters(simulation_size=(1, 1, H, W), device=dev, tau=0.6, contact_angle=None)
    st = S.random_state((H, W), 13, dev, obstacles=True, io_cells=True)
    flags = torch.ones_like(st["flags"])
    flags[:, :, 3:9, 700:1500] = st["flags"][:, :, 3:9, 700:1500]     # a patch of obstacles / inlets in a FLUID sea
    fs = jrl_step.FusedStep(P, operator="bgk")
    kw = dict(phi_obs=st["phi"], inflow_density=0.5, force=st["force"], inflow_vel=torch.tensor([0.01, 0.02], device=dev))

    def run(use_mask, fl):
        fs.use_tile_mask = use_mask
        return [t.cpu().numpy() for t in fs.step(st["f"], fl, st["rho"], st["vel"], force_vec=8, **kw)]
    a, b = run(True, flags), run(False, flags)
    assert fs._masks, "the mask was not built"
    for x, y in zip(a, b):
        assert np.array_equal(x, y, equal_nan=True)
    flags[:, :, 15, 2000:2100] = 2                                      # in-place edit: version counter moves
    a, b = run(True, flags), run(False, flags)
    for x, y in zip(a, b):
        assert np.array_equal(x, y, equal_nan=True)
    ref = [t.cpu().numpy() for t in fs.step(st["f"], flags, st["rho"], st["vel"], force_vec=4, **kw)]
    for x, y in zip(a, ref):
        assert np.array_equal(x, y, equal_nan=True)


@pytest.mark.parametrize("res,vec", [((40, 64), 0), ((21, 19), 0), ((24, 2048), 8), ((24, 2048), 4), ((24, 1040), 8)])
def test_step_sums(cuda_api, dev, res, vec):
    check_sums(cuda_api, dev, res=res, force_vec=vec)


def test_strip_flux_and_momentum_exchange(cuda_api, dev):
    import jrl_solver
    import lbm_oracle as O
    g = torch.Generator().manual_seed(3)
    C = torch.rand((1, 1, 40, 24), generator=g)
    mesh = cuda_api.create_2d_meshgrid_tensor((1, 1, 40, 24), dtype=torch.float32) + 0.5
    out = jrl_solver.strip_flux(C.to(dev), mesh.to(dev), 9, 7).cpu().numpy()
    assert abs(out[0] - O.strip_flux(C.numpy(), mesh.numpy(), 9, 7)) <= 1e-5 * abs(out[0])
    assert abs(out[1] - O.strip_sum(C.numpy()[0, 0], 9, 7)) <= 1e-5 * abs(out[1])
    st = S.random_state((33, 47), 5, dev)
    F = jrl_solver.momentum_exchange(st["f"], st["flags"]).cpu().numpy()
    Fo = O.momentum_exchange(st["f"].cpu().numpy(), st["flags"].cpu().numpy())
    assert np.allclose(F, np.asarray(Fo), rtol=1e-10, atol=1e-10)


def test_error_behaviour(cuda_api, dev):
    """The reference checks nothing; the drop-in raises RuntimeError and never touches memory."""
    P = cuda_api.SimulationParameters(simulation_size=(1, 1, 16, 16), device=dev)
    R = cuda_api.SimulationRunner(P)
    prop = R.create_propagation()
    f = torch.zeros((1, 9, 16, 16), device=dev)
    flags = torch.ones((1, 1, 16, 16), dtype=torch.uint8, device=dev)
    with pytest.raises(RuntimeError, match="no CPU path"):
        prop.propagation(f=f.cpu(), flags=flags.cpu())
    with pytest.raises(RuntimeError, match="dtype"):
        prop.propagation(f=f.double(), flags=flags)
    with pytest.raises(RuntimeError, match="contiguous"):
        prop.propagation(f=f.permute(0, 1, 3, 2), flags=flags)
    with pytest.raises(RuntimeError, match="KBC Type"):
        R.create_collision_KBC().collision(dx=1, dt=1, f=f, rho=f[:, :1].contiguous(), vel=f[:, :2].contiguous(),
                                           flags=flags, force=None, KBC_type=7)


@pytest.mark.parametrize("geom", ["plane", "45degree"])
def test_device_tip_motion_geometries(cuda_api, oracle_api, dev, geom):
    """batch_validation.py's other tips (infinite plane, 45-degree cone) on the device-side tip motion and the walk loop."""
    check_device_tip(cuda_api, dev, is_negative=True, fused=True, oracle=oracle_api, geom=geom, exact=True)


@pytest.mark.parametrize("geom", ["plane", "45degree"])
def test_fluid_walk_geometries(cuda_api, oracle_api, dev, geom):
    check_fluid_walk(cuda_api, dev, 2e-5, oracle=oracle_api, geom=geom, exact=True)
