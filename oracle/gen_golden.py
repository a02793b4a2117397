"""Test infrastructure (oracle/): write tests/golden/*.npz from the REFERENCE itself.

Runs every case of tests/golden_cases.py on the unmodified reference Python package
(/root/reference/src/LBM, imported on the CPU through oracle/ref_cpu.py, whose stand-in
extension modules execute the reference's own kernel bodies compiled by oracle/build_ref.sh).
Only runs where /root/reference exists; the fixtures it writes are committed.

    python oracle/gen_golden.py [--only NAME ...] [--check]

--check also runs the numpy oracle on each case and prints relative-L2 differences.
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm((a - b).ravel()) / max(1e-30, np.linalg.norm(b.ravel())))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*")
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    import torch
    import ref_cpu
    ref_cpu.import_reference()
    # jrl-lbm_b200/ must NOT be on sys.path here: its `src` package (the drop-in) would shadow
    # the reference's `src` namespace package.  Load the scenario builders by file path.
    import importlib.util
    spec = importlib.util.spec_from_file_location("jrl_scenarios", os.path.join(ROOT, "tests", "jrl_scenarios.py"))
    S = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(S)
    from golden_cases import CASES
    ref_api = S.api_from_src()
    if args.check:
        import oracle_api
        orc_api = oracle_api.api()
    outdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    cpu = torch.device("cpu")
    for name, (fn, kw) in CASES.items():
        if args.only and name not in args.only:
            continue
        t0 = time.time()
        ref = getattr(S, fn)(ref_api, cpu, **kw)
        ref = {k: v for k, v in ref.items() if v is not None}
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **ref)
        msg = f"{name}: reference {time.time() - t0:.1f}s"
        if args.check:
            t0 = time.time()
            orc = getattr(S, fn)(orc_api, cpu, **kw)
            worst = max(((rel_l2(orc[k], ref[k]), k) for k in ref if isinstance(ref[k], np.ndarray)
                         and ref[k].dtype.kind == "f"), default=(0, ""))
            exact = all(np.array_equal(orc[k], ref[k]) for k in ref if isinstance(ref[k], np.ndarray)
                        and ref[k].dtype.kind != "f")
            msg += f"; oracle {time.time() - t0:.1f}s worst rel-L2 {worst[0]:.3e} ({worst[1]}) ints exact={exact}"
        print(msg, flush=True)


if __name__ == "__main__":
    main()
