"""Compile the reference's four CUDA extensions from baseline/_ref/ (tools/make_baseline_ref.sh) for sm_100,
with the reference's own flags (with_cuda=True, --use_fast_math; LBM_propagation_2d.py:11-20 and three siblings),
into baseline/_ref/ext/<name>/<name>.so.  tools/bench_ref_cuda.py then hands these modules to the reference's
`torch.utils.cpp_extension.load(...)` calls on the GPU box instead of re-running the ~20-minute JIT there.

Not product code: baseline/_ref/ is git-ignored and holds the reference only.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
EXT = {
    "propagation": "LBM_propagation/cuda/propagation",
    "collision_2d": "LBM_collision/cuda/collision_2d",
    "macro_compute": "LBM_macro_compute/cuda/macro_compute",
    "solid_fluid_coupling": "LBM_solid_fluid_coupling/cuda/solid_fluid_coupling",
}


def build_one(name):
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"
    from torch.utils.cpp_extension import load
    stem = os.path.join(REF, "src", "LBM", EXT[name])
    out = os.path.join(REF, "ext", name)
    os.makedirs(out, exist_ok=True)
    load(name=name, sources=[stem + ".cpp", stem + ".cu"], verbose=True, with_cuda=True,
         extra_cuda_cflags=["--use_fast_math"], build_directory=out, is_python_module=False)
    print(f"prebuild_ref_ext: {name} -> {out}/{name}.so", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        build_one(sys.argv[1])
    else:
        procs = [subprocess.Popen([sys.executable, __file__, n]) for n in EXT]
        rc = [p.wait() for p in procs]
        if any(rc):
            sys.exit(f"prebuild_ref_ext: failed {dict(zip(EXT, rc))}")
