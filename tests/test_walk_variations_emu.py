"""The two walk loops below the ABI (jrl_conc_walk_f32, jrl_fluid_walk_f32: device-side tip motion + fused step + records,
no interpreter in the loop; SURVEY 8f ranks 1 + 2) against the ORACLE on parameter sets other than the one the named cases
use: other grid extents, tip radii, travel speeds (several cells crossed / none crossed), Rg, snapshot spacing, gravity
sign, both flag layouts.  Host-emulated kernels; floats <= 1e-5 relative L2 (util.field_tol), flags / masks / step counts exact."""
import pytest
import torch

import jrl_scenarios as S
from fused_cases import against_oracle, tip_snapshots

DEV = torch.device("cpu")

CONC = {
    "wide_fast": dict(res=(56, 40), balance_steps=4, walk_steps=50, radius_obs=3.50000000000001, inflow_height=10.50000000000001,
                      Rg=4, vel_obs_conv=0.06, is_negative=True, clamp=True, x_interval=0.25),
    "small_slow": dict(res=(40, 24), balance_steps=6, walk_steps=30, radius_obs=1.50000000000001, inflow_height=6.50000000000001,
                       Rg=2, vel_obs_conv=0.004, is_negative=True, clamp=True, x_interval=0.05),
    "positive_layout": dict(res=(44, 30), balance_steps=3, walk_steps=35, radius_obs=2.50000000000001,
                            inflow_height=7.50000000000001, Rg=3, vel_obs_conv=0.035, is_negative=False, clamp=False, x_interval=0.1),
    "plane": dict(res=(48, 28), balance_steps=4, walk_steps=40, radius_obs=2.50000000000001, inflow_height=9.50000000000001,
                  Rg=200, vel_obs_conv=0.02, is_negative=True, clamp=True, x_interval=0.2),
}


@pytest.mark.parametrize("name", list(CONC))
def test_conc_walk_variations_match_oracle(emu_api, oracle_api, name):
    kw = dict(CONC[name])
    n_snaps = int(kw["vel_obs_conv"] * kw["walk_steps"] / kw["x_interval"]) + 3
    kw["vel_snaps"] = tip_snapshots(emu_api, DEV, kw["res"], n=n_snaps, r0=kw["radius_obs"], inflow_height=kw["inflow_height"],
                                    x_interval=kw["x_interval"])
    run = S.jrl_conc(emu_api, DEV, device_tip="run", fused=True, **kw)
    against_oracle(oracle_api, "jrl_conc", kw, run, "jrl_conc_walk_" + name)


FLUID = {
    "wide_fast": dict(res=(56, 40), steps=60, radius_obs=3.50000000000001, inflow_height=10.50000000000001, Rg=4, vel_obs=0.04,
                      gravity_strength=2e-5),
    "small_slow_up": dict(res=(40, 24), steps=50, radius_obs=1.50000000000001, inflow_height=6.50000000000001, Rg=2, vel_obs=0.005,
                          gravity_strength=-1e-5),
    "plane_nog": dict(res=(48, 28), steps=40, radius_obs=2.50000000000001, inflow_height=9.50000000000001, Rg=200, vel_obs=0.02,
                      gravity_strength=0.0),
}


@pytest.mark.parametrize("name", list(FLUID))
def test_fluid_walk_variations_match_oracle(emu_api, oracle_api, name):
    kw = dict(FLUID[name])
    run = S.jrl_fluid(emu_api, DEV, device_tip="run", **kw)
    against_oracle(oracle_api, "jrl_fluid", kw, run, "a_" + name)
