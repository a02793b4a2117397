"""CPU: the C-ABI shared library (jrl-lbm_b200/libjrl_lbm.so, built by nvcc for sm_100a) loads
without a GPU and exports every symbol include/jrl_lbm.h declares; the ctypes prototype table
and descriptor structs of jrl_native.py match the header; compute entry points fail loudly
(never fall back) when there is no CUDA device."""
import ctypes
import os
import re

import pytest
import torch

from conftest import PKG, ROOT, have_cuda

HEADER = os.path.join(ROOT, "include", "jrl_lbm.h")
LIB = os.path.join(PKG, "libjrl_lbm.so")


def _header_text():
    with open(HEADER) as fh:
        txt = fh.read()
    return re.sub(r"/\*.*?\*/", "", txt, flags=re.S)  # drop comments


def header_functions():
    """{name: number of parameters} of every function the header declares."""
    out = {}
    for m in re.finditer(r"\b(?:int|long long|const char \*)\s*(jrl_\w+)\s*\(([^;{]*?)\)\s*;", _header_text(), flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


def header_struct_fields(name):
    m = re.search(r"typedef struct %s\s*\{(.*?)\}\s*%s;" % (name, name), _header_text(), flags=re.S)
    fields = []
    for decl in m.group(1).split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            fields.append(re.findall(r"\w+", re.sub(r"\[.*?\]", "", part))[-1])
    return fields


def test_header_declares_what_the_binding_binds():
    import jrl_native
    fns = header_functions()
    assert len(fns) >= 18
    assert set(fns) == set(jrl_native.PROTOTYPES), set(fns) ^ set(jrl_native.PROTOTYPES)
    for name, n in fns.items():
        assert len(jrl_native.PROTOTYPES[name]) == n, name


def test_descriptor_structs_match_the_header():
    import jrl_native
    assert [n for n, _ in jrl_native.StepDesc._fields_] == header_struct_fields("jrl_step_desc")
    assert [n for n, _ in jrl_native.CollideDesc._fields_] == header_struct_fields("jrl_collide_desc")
    assert [n for n, _ in jrl_native.TipDesc._fields_] == header_struct_fields("jrl_tip_desc")
    assert [n for n, _ in jrl_native.WalkDesc._fields_] == header_struct_fields("jrl_walk_desc")
    assert [n for n, _ in jrl_native.TipFluidDesc._fields_] == header_struct_fields("jrl_tip_fluid_desc")
    assert [n for n, _ in jrl_native.FluidWalkDesc._fields_] == header_struct_fields("jrl_fluid_walk_desc")


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "build it: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(LIB)
    for name in header_functions():
        assert hasattr(lib, name), f"libjrl_lbm.so lacks {name}"
    lib.jrl_version.restype = ctypes.c_int
    m = re.search(r"#define JRL_LBM_VERSION (\d+)", open(HEADER).read())
    assert lib.jrl_version() == int(m.group(1))


def test_library_is_sm100a_only():
    """The product library carries sm_100a SASS and nothing else (no multi-arch fat binary)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not installed")
    out = subprocess.run([cuobjdump, "--list-elf", LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


@pytest.mark.skipif(have_cuda(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    """Product code path on a machine without CUDA: the Python layer refuses CPU tensors and the
    C entry point itself returns JRL_ERR_CUDA -- nothing computes on the host."""
    import jrl_native
    saved = (jrl_native._lib, jrl_native._require_cuda)  # the session-scoped emu_api fixture may be live
    from host_emu import install as emu_install
    jrl_native._lib, jrl_native._require_cuda = None, emu_install._saved.get("req", jrl_native._require_cuda)
    try:
        jrl_native.load()
        import jrl_scenarios
        api = jrl_scenarios.api_from_src()
        P = api.SimulationParameters(simulation_size=(1, 1, 8, 8), device=torch.device("cpu"))
        prop = api.SimulationRunner(P).create_propagation()
        f = torch.zeros((1, 9, 8, 8))
        flags = torch.ones((1, 1, 8, 8), dtype=torch.uint8)
        with pytest.raises(RuntimeError, match="no CPU path"):
            prop.propagation(f=f, flags=flags)
        lib = ctypes.CDLL(LIB)
        lib.jrl_feq_f32.restype = ctypes.c_int
        lib.jrl_feq_f32.argtypes = jrl_native.PROTOTYPES["jrl_feq_f32"]
        lib.jrl_last_error.restype = ctypes.c_char_p
        buf = (ctypes.c_float * (11 * 64))()
        p = ctypes.addressof(buf)
        rc = lib.jrl_feq_f32(p, p + 256, None, p + 768, 1.0, 1.0, 1, 8, 8, None)
        assert rc == -2, rc  # JRL_ERR_CUDA
        assert lib.jrl_last_error()
    finally:
        jrl_native._lib, jrl_native._require_cuda = saved
