"""TEST INFRASTRUCTURE: run one of the reference's demo scripts BYTE-FOR-BYTE UNMODIFIED and capture its state.

The same runner serves both sides of the parity tests:
  * oracle/gen_golden_demos.py runs the script on the reference's own `src.LBM` (imported on the CPU through
    oracle/ref_cpu.py) and writes tests/golden/demo_*.npz;
  * tests/test_reference_demos_unchanged.py runs the SAME FILE with jrl-lbm_b200/ first on sys.path, so that
    `from src.LBM... import ...` resolves to the drop-in (host-emulated kernels here, libjrl_lbm.so on the B200).

How: the script is copied into a scratch `<workdir>/demo/` (the demos write next to `__file__` and the reference tree
is read-only; the copy's sha256 is checked against the source), executed with `runpy.run_path(..., "__main__")`
under its own argv from that directory (its `sys.path.append("../")` then adds nothing), and a `sys.setprofile`
hook copies `main()`'s local variables when it returns -- and at chosen step counts, detected as the N-th call of
`SimulationRunner.step()` from `main` -- so nothing in the script has to change to export f / rho / vel / C / h /
flags.  Text files the script wrote (records, balance) are read back.
"""
from __future__ import annotations

import contextlib
import hashlib
import io
import os
import runpy
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
STANDINS = os.path.join(HERE, "standins")


def demo_source_dir():
    """Where the reference's demo scripts can be read from: the reference tree in the build container, or the
    git-ignored copy tools/make_baseline_ref.sh leaves in baseline/_ref/ (which travels to the GPU box)."""
    for d in (os.environ.get("JRL_REFERENCE_DEMO_DIR"), "/root/reference/demo",
              os.path.join(os.path.dirname(HERE), "baseline", "_ref", "demo")):
        if d and os.path.isfile(os.path.join(d, "demo_2d_LBM.py")):
            return d
    return None


def ensure_standins():
    """imageio / h5py stand-ins (tests/standins/) only when the real modules are missing."""
    for mod in ("imageio", "h5py"):
        try:
            __import__(mod)
        except ImportError:
            if STANDINS not in sys.path:
                sys.path.append(STANDINS)
            __import__(mod)


def _sha(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def _export(v):
    """A main() local -> something comparable: tensors -> numpy copies; scalars kept; short lists recursed."""
    import torch
    if isinstance(v, torch.Tensor):
        return v.detach().to("cpu").numpy().copy()
    if isinstance(v, (bool, int, float, str)):
        return v
    if isinstance(v, (list, tuple)) and len(v) <= 16:
        out = [_export(x) for x in v]
        return out if all(x is not None for x in out) else None
    if isinstance(v, np.ndarray):
        return v.copy()
    return None


def _export_locals(loc, only=None):
    out = {}
    for k, v in loc.items():
        if only is not None and k not in only:
            continue
        e = _export(v)
        if e is not None:
            out[k] = e
    return out


def run_demo(script, argv, workdir, snapshots=(), snapshot_vars=None, quiet=True):
    """Run `<demo_source_dir>/<script>` unmodified.  Returns dict(final=locals of main at return,
    snaps={n: locals at the n-th SimulationRunner.step()}, workdir=<workdir>/demo, sha256=...)."""
    src_dir = demo_source_dir()
    if src_dir is None:
        raise FileNotFoundError("reference demo scripts not found (neither /root/reference/demo nor baseline/_ref/demo)")
    ensure_standins()
    demo_dir = os.path.join(workdir, "demo")
    os.makedirs(demo_dir, exist_ok=True)
    src = os.path.join(src_dir, script)
    dst = os.path.join(demo_dir, script)
    shutil.copyfile(src, dst)
    sha = _sha(src)
    assert _sha(dst) == sha
    snapshots = set(int(s) for s in snapshots)
    state = {"final": None, "snaps": {}, "nstep": 0}

    def hook(frame, event, arg):
        code = frame.f_code
        if event == "call" and code.co_name == "step" and snapshots:
            back = frame.f_back
            if back is not None and back.f_code.co_name == "main" and back.f_code.co_filename == dst:
                state["nstep"] += 1
                if state["nstep"] in snapshots:
                    state["snaps"][state["nstep"]] = _export_locals(back.f_locals, snapshot_vars)
        elif event == "return" and code.co_name == "main" and code.co_filename == dst:
            state["final"] = _export_locals(frame.f_locals)

    old_argv, old_cwd = sys.argv, os.getcwd()
    old_env = os.environ.get("TQDM_DISABLE")
    os.environ["TQDM_DISABLE"] = "1"
    sys.argv = [dst] + [str(a) for a in argv]
    os.chdir(demo_dir)
    sink = io.StringIO()
    try:
        with (contextlib.redirect_stdout(sink) if quiet else contextlib.nullcontext()), \
                (contextlib.redirect_stderr(io.StringIO()) if quiet else contextlib.nullcontext()):
            sys.setprofile(hook)
            try:
                runpy.run_path(dst, run_name="__main__")
            except SystemExit as e:  # the JRL demos call exit(0) on "already done" / "fluid field missing"
                state["exit"] = e.code
            except Exception as e:  # noqa: BLE001 -- e.g. the stage-B demos index one velocity snapshot past the end
                # of the stage-A cache in their last steps (demo_2d_LBM_JRL_C_cylinder.py:295-302 with
                # demo_2d_LBM_JRL_fluid_cylinder.py:239-250): the reference itself ends that way; main()'s locals at
                # the raise are still captured by the hook, and the drop-in must end at the same step.
                state["error"] = f"{type(e).__name__}: {e}"
            finally:
                sys.setprofile(None)
    finally:
        os.chdir(old_cwd)
        sys.argv = old_argv
        if old_env is None:
            os.environ.pop("TQDM_DISABLE", None)
        else:
            os.environ["TQDM_DISABLE"] = old_env
    return dict(final=state["final"], snaps=state["snaps"], workdir=demo_dir, sha256=sha, stdout=sink.getvalue(),
                exit=state.get("exit"), error=state.get("error"))


def read_series(path):
    """`record_*.txt` / `balance.txt`: header line, then whitespace-separated floats -> float64 [n, cols]."""
    rows = []
    with open(path) as f:
        next(f)
        for line in f:
            if line.strip():
                rows.append([float(x) for x in line.split()])
    return np.asarray(rows, dtype=np.float64)


def find_files(root, suffix):
    out = []
    for d, _, files in os.walk(root):
        out += [os.path.join(d, f) for f in files if f.endswith(suffix)]
    return sorted(out)


# ----------------------------------------------------------------------------------------------------------
# case-level driver shared by oracle/gen_golden_demos.py and tests/test_reference_demos_unchanged.py
# ----------------------------------------------------------------------------------------------------------
def apply_view(arr, view):
    a = np.asarray(arr)
    if view == "full":
        return a
    if view == "stride2":
        return a[..., ::2, ::2]
    if view == "sum":
        return a.astype(np.float64).sum(axis=(-2, -1))
    if view == "snapsum":
        return a.astype(np.float64).sum(axis=(-2, -1))
    if isinstance(view, tuple) and view[0] == "crop":
        _, y0, y1, x0, x1 = view
        return a[..., y0:y1, x0:x1]
    if isinstance(view, tuple) and view[0] == "snaps":
        return a[[i for i in view[1:]]]
    raise ValueError(view)


def view_name(view):
    return view if isinstance(view, str) else "_".join(str(x) for x in view)


def extract(result, case, cpu=False):
    """Flatten a run_demo() result into {"<tag>.<variable>.<view>": array} following case["keep"]."""
    out = {}
    cpu_only = cpu and case.get("cpu_steps", "all") != "all"
    for tag, keep in case["keep"].items():
        if tag == "final":
            if cpu_only:
                continue  # the CPU suite stops at the snapshot count; its final state IS the snapshot
            loc = result["final"]
        elif cpu_only and tag == case["cpu_steps"]:
            loc = result["final"]
        else:
            loc = result["snaps"].get(tag)
        if loc is None:
            raise KeyError(f"no captured state for tag {tag}")
        for var, views in keep.items():
            if var not in loc:
                raise KeyError(f"main() has no local `{var}` at {tag}")
            for view in (views if isinstance(views, list) else [views]):
                out[f"{tag}.{var}.{view_name(view)}"] = np.asarray(apply_view(loc[var], view))
    return out


def run_case(name, cases, workdir, cpu=False):
    """Run case `name` (after the stage-A case it needs, in the same scratch demo dir) -> flat dict of views,
    plus the record / balance series and the balanced C, h cache for stage-B cases."""
    import shutil as _sh
    from demo_cases import fluid_cache_name, stage_prefix
    case = cases[name]
    demo_dir = os.path.join(workdir, "demo")
    if case.get("needs"):
        a_case = cases[case["needs"]]
        src = os.path.join(demo_dir, stage_prefix(a_case), fluid_cache_name())
        if not os.path.exists(src):
            ra = run_demo(a_case["script"], a_case["argv"], workdir)
            assert ra["final"] is not None, f"stage A {case['needs']} did not finish: {ra['stdout'][-500:]}"
        dst = os.path.join(demo_dir, stage_prefix(case), fluid_cache_name())
        if src != dst:
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            _sh.copyfile(src, dst)
    cpu_only = cpu and case.get("cpu_steps", "all") != "all"
    argv = case["argv_cpu"] if cpu_only else case["argv"]
    snaps = () if cpu_only else case["snapshots"]
    res = run_demo(case["script"], argv, workdir, snapshots=snaps)
    if res["final"] is None or "f" not in res["final"] and "h" not in res["final"]:
        raise RuntimeError(f"{name}: main() did not run to the end (exit={res['exit']}): {res['stdout'][-800:]}")
    out = extract(res, case, cpu=cpu)
    if case.get("stage") == "B":
        pre = os.path.join(demo_dir, stage_prefix(case))
        rec = [p for p in find_files(pre, ".txt") if os.path.basename(p).startswith("record_")]
        bal = [p for p in find_files(pre, ".txt") if os.path.basename(p) == "balance.txt"]
        assert len(rec) == 1 and len(bal) == 1, (rec, bal)
        out["records"] = read_series(rec[0])
        out["balance"] = read_series(bal[0])
        import h5py
        hb = h5py.File(os.path.join(pre, "C_balance.h5"))
        out["balance.C"], out["balance.h"] = np.asarray(hb["C"]), np.asarray(hb["h"])
        hb.close()
    out["error"] = np.asarray(res["error"] or "")
    out["last_step"] = np.asarray(int(res["final"].get("step", -1)))
    out["script_sha256"] = np.frombuffer(bytes.fromhex(res["sha256"]), dtype=np.uint8)
    return out
