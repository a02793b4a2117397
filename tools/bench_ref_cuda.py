"""Reference-as-run on a B200 (SURVEY 8c "Oracle B", 8d "reference timing beside it").

The reference hard-requires CUDA and is built with --use_fast_math (readme.md:7, LBM_propagation_2d.py:11-20), so the
host-compiled IEEE oracle is not what its users run.  This tool runs the reference ITSELF -- baseline/_ref/ =
tools/make_baseline_ref.sh's copy (seven `.type()` -> `.scalar_type()` edits, nothing else), its four extensions
either prebuilt for sm_100 by tools/prebuild_ref_ext.py or JIT-built on the spot -- on cuda:0 and reports

  (i)  the time per step of the reference's own loop bodies (tests/jrl_scenarios.py restates them call for call) at the
       native grids (130 x 130, 800 x 400, 128 x 64) and at 8192 x 8192, by CUDA-synchronised wall clock over the
       difference of two runs with different step counts (set-up cancels), next to this library through the SAME
       solver-object calls and through its fused step;
  (ii) the difference between this library and that fast-math build on every case of tests/golden_cases.py
       (per-operator single calls and the C1-C5 scenarios) and on the unmodified demo scripts C1-C3: relative L2 and
       the count of differing bits, integers / masks exact.

Two processes (both packages are called `src`):   --role ref  writes gpurun_out/refcuda/*.npz + timings,
                                                  --role ours compares and times the drop-in,
no --role runs both and prints / writes the summary JSON (gpurun_out/refcuda/summary.json).
The reference outputs are also what tests/golden/refcuda_*.npz are copied from (tests/test_gpu_vs_refcuda.py).
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
OUT = os.path.join(ROOT, "gpurun_out", "refcuda")
sys.path.append(os.path.join(ROOT, "tests"))

# (name, scenario function, kwargs without `steps`, (n1, n2) step counts) -- MLUPS = cells * (n2 - n1) / (t2 - t1)
JRL = dict(radius_obs=6.50000000000001, inflow_height=20.50000000000001, Rg=4)
TIMED = [
    ("c1_bgk_130", "shear_layer", dict(res=(130, 130), operator="bgk"), (50, 350)),
    ("c2_mrt_130", "shear_layer", dict(res=(130, 130), operator="mrt", kbc_type=133), (50, 350)),
    ("c2_kbc_a_130", "shear_layer", dict(res=(130, 130), operator="kbc", kbc_type=133), (50, 250)),
    ("c3_cylinder_800x400", "cylinder", dict(res=(800, 400), D=20), (30, 230)),
    ("c5_stageA_128x64", "jrl_fluid", dict(res=(128, 64), vel_obs=0.002, gravity_strength=2e-5, **JRL), (30, 230)),
    ("c4_conc_128x64", "jrl_conc", dict(res=(128, 64), balance_steps=0, vel_obs_conv=0.01, **JRL), (30, 230)),
    ("c1_bgk_8192", "shear_layer", dict(res=(8192, 8192), operator="bgk"), (3, 23)),
    ("c2_mrt_8192", "shear_layer", dict(res=(8192, 8192), operator="mrt", kbc_type=133), (3, 13)),
]
DEMOS = ["c1_bgk", "c2_mrt", "c3_cylinder"]


def setup_reference():
    """Import the reference's src.LBM on CUDA.  Prebuilt extension modules are handed to its `load(...)` calls."""
    import torch.utils.cpp_extension as cpp_ext
    real_load = cpp_ext.load

    def load(name, sources, **kw):
        so = os.path.join(REF, "ext", name, name + ".so")
        if os.path.exists(so):
            spec = importlib.util.spec_from_file_location(name, so)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            print(f"[ref] {name}: prebuilt {so}", flush=True)
            return mod
        t0 = time.time()
        mod = real_load(name, sources, **kw)   # cwd = baseline/_ref/demo, as the reference's relative paths need
        print(f"[ref] {name}: JIT-built in {time.time() - t0:.0f} s", flush=True)
        return mod

    cpp_ext.load = load
    os.chdir(os.path.join(REF, "demo"))
    sys.path.insert(0, REF)
    import demo_runner
    demo_runner.ensure_standins()
    import src.LBM.simulation  # noqa: F401
    assert os.path.realpath(sys.modules["src"].__path__[0] if hasattr(sys.modules["src"], "__path__") else "").startswith(
        os.path.realpath(REF)), "the reference package was shadowed"


def setup_ours():
    sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200"))
    import jrl_native
    jrl_native.load()
    import src.LBM.simulation  # noqa: F401


def timed(S, api, dev, fn, kw, n, **extra):
    import torch
    key = "walk_steps" if fn == "jrl_conc" else "steps"
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    getattr(S, fn)(api, dev, **kw, **{key: n}, **extra)
    torch.cuda.synchronize()
    return time.perf_counter() - t0


def time_cases(S, api, dev, only, variants):
    rows = []
    for name, fn, kw, (n1, n2) in TIMED:
        if only and name not in only:
            continue
        cells = kw["res"][0] * kw["res"][1]
        for label, extra in variants(fn, kw):
            try:
                timed(S, api, dev, fn, kw, 2, **extra)  # warm-up (allocator, lazy module loads)
                t1 = timed(S, api, dev, fn, kw, n1, **extra)
                t2 = timed(S, api, dev, fn, kw, n2, **extra)
                us = (t2 - t1) / (n2 - n1) * 1e6
                rows.append(dict(case=name, path=label, us_per_step=round(us, 2), mlups=round(cells / us, 1),
                                 steps=(n1, n2)))
            except Exception as e:  # noqa: BLE001
                rows.append(dict(case=name, path=label, error=f"{type(e).__name__}: {e}"[:300]))
            print(json.dumps(rows[-1]), flush=True)
    return rows


def bitdiff(a, b):
    import numpy as np
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    if a.dtype.kind == "f":
        return int((a.view(np.uint32 if a.dtype.itemsize == 4 else np.uint64)
                    != b.view(np.uint32 if b.dtype.itemsize == 4 else np.uint64)).sum())
    return int((a != b).sum())


def rel_l2(a, b):
    import numpy as np
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    m = np.isfinite(a) & np.isfinite(b)
    return float(np.linalg.norm(a[m] - b[m]) / max(1e-30, np.linalg.norm(b[m])))


def role_ref(args):
    import numpy as np
    import torch
    setup_reference()
    import jrl_scenarios as S
    import demo_runner as D
    from demo_cases import DEMO_CASES
    from golden_cases import CASES
    os.makedirs(OUT, exist_ok=True)
    api, dev = S.api_from_src(), torch.device("cuda:0")
    if not args.skip_goldens:
        for name, (fn, kw) in ({} if args.no_scenarios else CASES).items():
            out = getattr(S, fn)(api, dev, **kw)
            np.savez_compressed(os.path.join(OUT, f"refcuda_{name}.npz"), **{k: v for k, v in out.items() if v is not None})
            print(f"[ref] scenario {name} done", flush=True)
        import tempfile
        work = tempfile.mkdtemp(prefix="jrl_refcuda_ref_")
        demo_s = {}
        for name in args.demos:
            t0 = time.time()
            flat = D.run_case(name, DEMO_CASES, work, cpu=False)
            torch.cuda.synchronize()
            demo_s[name] = round(time.time() - t0, 2)
            np.savez_compressed(os.path.join(OUT, f"demo_refcuda_{name}.npz"), **flat)
            print(f"[ref] demo {name}: {demo_s[name]} s", flush=True)
            with open(os.path.join(OUT, "demo_seconds_ref.json"), "w") as f:
                json.dump(demo_s, f, indent=1)
    if args.no_timing:
        return
    rows = time_cases(S, api, dev, args.only, lambda fn, kw: [("reference (torch-CUDA, as shipped)", {})])
    with open(os.path.join(OUT, "timing_ref.json"), "w") as f:
        json.dump(rows, f, indent=1)


def role_ours(args):
    import numpy as np
    import torch
    setup_ours()
    import jrl_scenarios as S
    import demo_runner as D
    from demo_cases import DEMO_CASES
    from golden_cases import CASES
    api, dev = S.api_from_src(), torch.device("cuda:0")
    diffs, demo_s = [], {}
    if not args.skip_goldens:
        for name, (fn, kw) in ({} if args.no_scenarios else CASES).items():
            path = os.path.join(OUT, f"refcuda_{name}.npz")
            if not os.path.exists(path):
                continue
            with np.load(path) as z:
                ref = {k: z[k] for k in z.files}
            got = getattr(S, fn)(api, dev, **kw)
            for k, w in ref.items():
                g = np.asarray(got[k])
                if w.ndim == 0:
                    continue
                row = dict(case=name, key=k, n=int(w.size), bits_differ=bitdiff(g, w))
                if w.dtype.kind == "f":
                    row["rel_l2"] = rel_l2(g, w)
                diffs.append(row)
        import tempfile
        work = tempfile.mkdtemp(prefix="jrl_refcuda_ours_")
        for name in args.demos:
            path = os.path.join(OUT, f"demo_refcuda_{name}.npz")
            if not os.path.exists(path):
                continue
            with np.load(path) as z:
                ref = {k: z[k] for k in z.files}
            t0 = time.time()
            got = D.run_case(name, DEMO_CASES, work, cpu=False)
            torch.cuda.synchronize()
            demo_s[name] = round(time.time() - t0, 2)
            with open(os.path.join(OUT, "demo_seconds_ours.json"), "w") as f:
                json.dump(demo_s, f, indent=1)
            for k, w in ref.items():
                if w.dtype.kind in "US" or w.ndim == 0:
                    continue
                row = dict(case="demo_" + name, key=k, n=int(w.size), bits_differ=bitdiff(got[k], w))
                if w.dtype.kind == "f":
                    row["rel_l2"] = rel_l2(got[k], w)
                diffs.append(row)
        for r in diffs:
            print(json.dumps(r), flush=True)

    def variants(fn, kw):
        v = [("this library, same solver-object calls", {})]
        if fn in ("shear_layer", "cylinder"):
            v.append(("this library, FusedStep.run", {"fused": "run"}))
        return v
    with open(os.path.join(OUT, "diffs.json"), "w") as f:
        json.dump(diffs, f, indent=1)
    if args.no_timing:
        return
    rows = time_cases(S, api, dev, args.only, variants)
    with open(os.path.join(OUT, "timing_ours.json"), "w") as f:
        json.dump(rows, f, indent=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--role", choices=["ref", "ours"])
    ap.add_argument("--only", nargs="*")
    ap.add_argument("--demos", nargs="*", default=DEMOS)
    ap.add_argument("--skip-goldens", action="store_true")
    ap.add_argument("--no-scenarios", action="store_true")
    ap.add_argument("--no-timing", action="store_true")
    args = ap.parse_args()
    if args.role == "ref":
        return role_ref(args)
    if args.role == "ours":
        return role_ours(args)
    os.makedirs(OUT, exist_ok=True)
    extra = sys.argv[1:]
    for role in ("ref", "ours"):
        with open(os.path.join(OUT, f"log_{role}.txt"), "w") as log:
            rc = subprocess.run([sys.executable, os.path.abspath(__file__), "--role", role] + extra, stdout=log,
                                stderr=subprocess.STDOUT).returncode
        print(f"role {role}: exit {rc}", flush=True)
    summary = {}
    for k in ("timing_ref", "timing_ours", "diffs", "demo_seconds_ref", "demo_seconds_ours"):
        p = os.path.join(OUT, k + ".json")
        if os.path.exists(p):
            with open(p) as f:
                summary[k] = json.load(f)
    worst = {}
    for r in summary.get("diffs", []):
        if "rel_l2" in r:
            w = worst.setdefault(r["case"], dict(rel_l2=0.0, key="", float_bits_differ=0, n=0))
            if r["rel_l2"] >= w["rel_l2"]:
                w["rel_l2"], w["key"] = r["rel_l2"], r["key"]
            w["float_bits_differ"] += r["bits_differ"]
            w["n"] += r["n"]
        else:
            worst.setdefault(r["case"], dict(rel_l2=0.0, key="", float_bits_differ=0, n=0)).setdefault("int_mismatch", 0)
            worst[r["case"]]["int_mismatch"] = worst[r["case"]].get("int_mismatch", 0) + r["bits_differ"]
    summary["worst_per_case"] = worst
    with open(os.path.join(OUT, "summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    print(json.dumps(worst, indent=1))


if __name__ == "__main__":
    main()
