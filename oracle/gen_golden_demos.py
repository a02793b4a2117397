"""Test infrastructure (oracle/): goldens from the reference's OWN demo scripts, run unmodified.

For every case of tests/demo_cases.py the script file under /root/reference/demo is executed byte-for-byte as
shipped (tests/demo_runner.py: runpy + a profile hook that copies main()'s locals) on the reference's own Python
package, imported on the CPU through oracle/ref_cpu.py (stand-in extension modules that execute the reference's
kernel bodies compiled by oracle/build_ref.sh; imageio / h5py stand-ins from tests/standins/).  The selected views
go to tests/golden/demo_<case>.npz, which are committed; /root/reference is not needed to run the tests.

    python oracle/gen_golden_demos.py [--only NAME ...]
"""
import argparse
import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*")
    args = ap.parse_args()
    import ref_cpu
    ref_cpu.import_reference()
    import demo_runner as D
    from demo_cases import DEMO_CASES
    outdir = os.path.join(ROOT, "tests", "golden")
    with tempfile.TemporaryDirectory(prefix="jrl_ref_demo_") as work:
        for name, case in DEMO_CASES.items():
            if args.only and name not in args.only:
                continue
            t0 = time.time()
            flat = D.run_case(name, DEMO_CASES, work, cpu=False)
            path = os.path.join(outdir, f"demo_{name}.npz")
            np.savez_compressed(path, **flat)
            bad = [k for k, v in flat.items() if v.dtype.kind == "f" and not np.isfinite(v).all()]
            print(f"{name}: {time.time() - t0:.1f}s, {len(flat)} arrays, {os.path.getsize(path) / 1024:.0f} KB"
                  + (f"  NON-FINITE: {bad}" if bad else ""), flush=True)


if __name__ == "__main__":
    main()
