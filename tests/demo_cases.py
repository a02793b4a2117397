"""Case table of the UNMODIFIED-reference-demo parity tests (SURVEY 8c sizes and step counts).

Shared by oracle/gen_golden_demos.py (runs each script on the reference itself, on the CPU, and writes
tests/golden/demo_<name>.npz) and tests/test_reference_demos_unchanged.py (runs the same script file on the
drop-in and compares).  A case:

    script     file name under the reference's demo/
    argv       its command line (native grids; the JRL runs use a large --Re / --vel_obs_real / --Pe so that the tip
               crosses the whole domain in ~2 000 steps instead of 11-14 million: total_steps = travel / vel_obs,
               vel_obs = Re * visc / radius_obs, demo_2d_LBM_JRL_fluid_cylinder.py:78,201)
    snapshots  step counts (N-th SimulationRunner.step() call) at which main()'s locals are captured as well
    keep       {variable of main(): view}; view = "full" | "stride2" | ("crop", y0, y1, x0, x1) | "sum"
    cpu_steps  what the CPU-only suite runs (host-emulated kernels are slow): "all" or a snapshot count, in which case
               argv_cpu replaces argv
    needs      name of the stage-A case whose velocity cache this stage-B case loads (batch_validation.py pairs)
"""

_C12_KEEP = {"density": "full", "vel": "full"}
_C12_FINAL = {"density": "full", "vel": "full", "f": "full", "flags": "full"}
_C3_VIEWS = {"density": ["stride2", ("crop", 140, 340, 130, 270), "sum"],
             "vel": ["stride2", ("crop", 140, 340, 130, 270), "sum"]}
_C3_FINAL = dict(_C3_VIEWS, f=[("crop", 170, 260, 160, 240), "sum"], flags=["full"], phi_obs=["stride2"])

# bounded JRL parameters (see module docstring): lattice tip speed 0.0513 (fluid) / 0.05 (concentration)
_RE, _VREAL, _PE = 2.0, 0.05, 1.95
_BAL = 0.000962          # --balance_time giving balance_step = 500 (demo_2d_LBM_JRL_C_cylinder.py:215)
_A_KEEP = {"f": ["full"], "density": ["full"], "vel": ["full"], "flags": ["full"], "phi_obs": ["full"],
           "vel_field": [("snaps", 0, 350, 700, -1), "snapsum"], "walks": ["full"], "gravity_strength": ["full"],
           "vel_obs": ["full"], "total_steps": ["full"]}
_B_KEEP = {"h": ["full"], "C": ["full"], "C_flags": ["full"], "C_vel": ["full"], "phi_obs": ["full"], "walks": ["full"],
           "tau_D": ["full"], "vel_obs_conv": ["full"], "total_steps": ["full"], "balance_step": ["full"]}


def _fluid(geom, g):
    return dict(script=f"demo_2d_LBM_JRL_fluid_{geom}.py",
                argv=["--res", 128, 64, "--Re", _RE, "--vel_obs_real", _VREAL, "--gravity_strength_real", g],
                snapshots=(), keep={"final": _A_KEEP}, cpu_steps="all", stage="A", geom=geom, g=g)


def _conc(geom, g, needs):
    return dict(script=f"demo_2d_LBM_JRL_C_{geom}.py",
                argv=["--res", 128, 64, "--Re", _RE, "--Pe", _PE, "--vel_obs_real", _VREAL, "--gravity_strength_real", g,
                      "--balance_time", _BAL, "--is_negative"],
                snapshots=(200,), keep={"final": _B_KEEP, 200: {"C": ["full"], "h": ["full"], "C_flags": ["full"]}},
                cpu_steps="all", stage="B", geom=geom, g=g, needs=needs)


DEMO_CASES = {
    # C1: demo_2d_LBM.py exactly as shipped (130 x 130, 1000 steps), state also captured after 100 steps
    "c1_bgk": dict(script="demo_2d_LBM.py", argv=["--res", 130, 130, "--total_steps", 1000], snapshots=(100,),
                   keep={100: _C12_KEEP, "final": _C12_FINAL}, cpu_steps=100,
                   argv_cpu=["--res", 130, 130, "--total_steps", 100]),
    # C2: demo_2d_LBM_KBC.py as shipped (it runs MRT)
    "c2_mrt": dict(script="demo_2d_LBM_KBC.py", argv=["--res", 130, 130, "--total_steps", 1000], snapshots=(100,),
                   keep={100: _C12_KEEP, "final": _C12_FINAL}, cpu_steps=100,
                   argv_cpu=["--res", 130, 130, "--total_steps", 100]),
    # C3: demo_2d_LBM_cylinder.py at its native 800 x 400 grid, 1000 of its 50 000 steps
    "c3_cylinder": dict(script="demo_2d_LBM_cylinder.py", argv=["--res", 800, 400, "--total_steps", 1000],
                        snapshots=(100,), keep={100: _C3_VIEWS, "final": _C3_FINAL}, cpu_steps=100,
                        argv_cpu=["--res", 800, 400, "--total_steps", 100]),
    # C5 = batch_validation.py:26-55: four stage-A -> stage-B pairs
    "a_45degree_g9": _fluid("45degree", -9.8),
    "b_45degree_g9": _conc("45degree", -9.8, "a_45degree_g9"),
    "a_45degree_g0": _fluid("45degree", 0),
    "b_45degree_g0": _conc("45degree", 0, "a_45degree_g0"),
    "a_cylinder_g9": _fluid("cylinder", -9.8),
    "b_cylinder_g9": _conc("cylinder", -9.8, "a_cylinder_g9"),   # = C4 at its native 128 x 64
    "a_InfinitePlane_g9": _fluid("InfinitePlane", -9.8),
    "b_InfinitePlane_g9": _conc("InfinitePlane", -9.8, "a_InfinitePlane_g9"),
}


def stage_prefix(case):
    """Directory (relative to the demo dir) where a JRL demo keeps its caches: fluid_cylinder writes `_g9/`, while
    C_cylinder reads `_cylinder_g9/` (batch_validation step 3 does not chain as shipped; the harness
    copies the file across, which is what a user has to do by hand)."""
    g = int(abs(case["g"]))
    if case["geom"] == "cylinder":
        return f"_g{g}" if case["stage"] == "A" else f"_cylinder_g{g}"
    return f"_{case['geom']}_g{g}"


def fluid_cache_name():
    return f"Re_{_RE}.h5"
