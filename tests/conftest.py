"""pytest configuration.

Markers: `gpu` = needs a CUDA device (run on the B200 box with `-m gpu`); everything else runs on
the CPU-only build container (`-m "not gpu"`).

Import paths: tests import the product as the reference's demos do (`src.LBM...`, with
jrl-lbm_b200/ on sys.path) and the checker from oracle/.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "jrl-lbm_b200")
ORACLE = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
EMU_LIB = os.path.join(ROOT, "tests", "host_emu", "libjrl_lbm_emu.so")
for p in (PKG, ORACLE, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def have_cuda():
    import torch
    return torch.cuda.is_available()


def pytest_collection_modifyitems(config, items):
    if have_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def build_emu():
    """tests/host_emu: the kernel sources of jrl-lbm_b200/csrc compiled for the host (test-only).
    JRL_EMU_LIB in the environment substitutes another build of the same sources (tools/sanitize_emu.sh: ASan + UBSan)."""
    if os.environ.get("JRL_EMU_LIB"):
        return os.environ["JRL_EMU_LIB"]
    srcs = [os.path.join(PKG, "csrc", f) for f in os.listdir(os.path.join(PKG, "csrc"))]
    srcs += [os.path.join(ROOT, "include", "jrl_lbm.h"), os.path.join(ROOT, "tests", "host_emu", "cuda_runtime.h"),
             os.path.join(ROOT, "tests", "host_emu", "emu_globals.cpp")]
    if not os.path.exists(EMU_LIB) or os.path.getmtime(EMU_LIB) < max(os.path.getmtime(s) for s in srcs):
        subprocess.run(["bash", os.path.join(ROOT, "tests", "host_emu", "build.sh")], check=True,
                       stdout=subprocess.DEVNULL)
    return EMU_LIB


@pytest.fixture(scope="session")
def emu_api():
    """The product's Python host layer (src.LBM) running on the host-compiled kernels: checks the
    host logic and the kernel logic where no GPU exists.  Never used by -m gpu tests."""
    if have_cuda():
        pytest.skip("host emulation is for the CPU-only container; on a GPU box the CUDA library is tested")
    from host_emu import install as emu_install
    emu_install.install(build_emu())
    import jrl_scenarios
    api = jrl_scenarios.api_from_src()
    yield api
    emu_install.uninstall()


@pytest.fixture(scope="session")
def cuda_api():
    """The product on the real library (libjrl_lbm.so, sm_100a).  Fails loudly if it is missing."""
    import jrl_native
    jrl_native.load()
    import jrl_scenarios
    return jrl_scenarios.api_from_src()


@pytest.fixture(scope="session")
def oracle_api():
    import oracle_api as OA
    return OA.api()
