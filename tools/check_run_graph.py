"""GPU check (not a pytest: needs >= 1024 steps): jrl_run_f32's CUDA-graph replay path (small grids, nsteps >= 1024) gives the
same bits as one jrl_step_f32 per step, and its time per step."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "jrl-lbm_b200")); sys.path.append(os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import jrl_native  # noqa: E402
jrl_native.load()
import jrl_scenarios as S  # noqa: E402

api, dev = S.api_from_src(), torch.device("cuda:0")
for fn, kw, n in (("shear_layer", dict(res=(130, 130), operator="bgk"), 1500),
                  ("shear_layer", dict(res=(130, 130), operator="mrt", kbc_type=133), 1301),
                  ("cylinder", dict(res=(800, 400), D=20), 1100)):
    a = getattr(S, fn)(api, dev, steps=n, fused=True, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    b = getattr(S, fn)(api, dev, steps=n, fused="run", **kw)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    c = getattr(S, fn)(api, dev, steps=11 * n, fused="run", **kw)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    same = all(np.array_equal(a[k], b[k], equal_nan=True) for k in ("f", "rho", "vel"))
    print("   mode", os.environ.get("JRL_RUN_GRAPH", "default"), "differing values:",
          {k: int((~np.isclose(a[k], b[k], rtol=0, atol=0, equal_nan=True)).sum()) for k in ("f", "rho", "vel")}, "max |diff| f: %.3e" % float(np.abs(a["f"] - b["f"]).max()))
    print(fn, kw.get("operator", "mrt+wall"), n, "steps: graph replay == per-step launches:", bool(same),
          " us/step (differenced): %.2f" % (((t2 - t1) - (t1 - t0)) / (10 * n) * 1e6), flush=True)
    assert same and np.isfinite(b["f"]).all()
