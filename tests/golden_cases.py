"""Case table shared by oracle/gen_golden.py (which runs each case on the REFERENCE, imported on
the CPU, and writes tests/golden/<name>.npz) and the parity tests (which run the same case on
the oracle and on the CUDA drop-in).  Each entry: name -> (jrl_scenarios function name, kwargs).
Sizes are chosen so one case takes seconds on the oracle and a fixture stays tens of KB."""

CASES = {
    # C1: demo_2d_LBM.py, periodic BGK (non-square grid to catch H/W swaps)
    "c1_shear_bgk": ("shear_layer", dict(res=(40, 32), steps=60, operator="bgk", snapshots=(10,))),
    # C2: demo_2d_LBM_KBC.py as shipped (MRT, KBC flag ignored) ...
    "c2_shear_mrt": ("shear_layer", dict(res=(40, 32), steps=60, operator="mrt", kbc_type=133)),
    # ... and as the config's wording asks (entropic KBC), all four variants
    "c2_shear_kbc_a": ("shear_layer", dict(res=(32, 32), steps=12, operator="kbc", kbc_type=133)),
    "c2_shear_kbc_b": ("shear_layer", dict(res=(32, 32), steps=12, operator="kbc", kbc_type=134)),
    "c2_shear_kbc_c": ("shear_layer", dict(res=(32, 32), steps=12, operator="kbc", kbc_type=137)),
    "c2_shear_kbc_d": ("shear_layer", dict(res=(32, 32), steps=12, operator="kbc", kbc_type=138)),
    # C3: demo_2d_LBM_cylinder.py scaled down (H=96, W=48, D=8)
    "c3_cylinder": ("cylinder", dict(res=(96, 48), D=8, steps=60, center=(24.0, 30.0), snapshots=(20,))),
    # C4/C5 stage A: axisymmetric moving-tip flow, with and without gravity; fast tip so it crosses cells
    "c5_jrl_fluid_g": ("jrl_fluid", dict(res=(48, 32), steps=80, radius_obs=2.50000000000001,
                                        inflow_height=8.50000000000001, Rg=3, vel_obs=0.02,
                                        gravity_strength=2e-5)),
    "c5_jrl_fluid_nog": ("jrl_fluid", dict(res=(48, 32), steps=40, radius_obs=2.50000000000001,
                                          inflow_height=8.50000000000001, Rg=3, vel_obs=0.02)),
    # C4/C5 stage B: concentration field, balance + walk phases, flux records
    "c4_jrl_conc": ("jrl_conc", dict(res=(48, 32), balance_steps=40, walk_steps=60, radius_obs=2.50000000000001,
                                    inflow_height=8.50000000000001, Rg=3, vel_obs_conv=0.03)),
    "c4_jrl_conc_positive": ("jrl_conc", dict(res=(40, 28), balance_steps=10, walk_steps=30,
                                             radius_obs=2.50000000000001, inflow_height=8.50000000000001, Rg=3,
                                             vel_obs_conv=0.03, is_negative=False, clamp=False)),
    # per-operator single calls on random states: plain, x-axis, y-axis symmetry; tensor contact angle
    "ops_plain": ("single_calls", dict(res=(20, 24), seed=1, axisym=0, contact_angle=1.2)),
    "ops_axis_x": ("single_calls", dict(res=(18, 14), seed=2, axisym=1)),
    "ops_axis_y": ("single_calls", dict(res=(14, 18), seed=3, axisym=2, contact_angle=0.5 * 3.141592653589793)),
    "ops_batch2": ("single_calls", dict(res=(12, 16), seed=4, axisym=1, B=2)),
}
