"""CPU: the product's Python host layer + kernel sources (compiled for the host, tests/host_emu)
against the golden fixtures written by the REFERENCE (oracle/gen_golden.py) and the numpy oracle.
The same cases run on the real CUDA library in test_gpu_parity.py."""
import numpy as np
import pytest
import torch

import jrl_scenarios as S
from golden_cases import CASES
from util import compare_outputs, load_golden


@pytest.mark.parametrize("name", list(CASES))
def test_case_matches_reference_golden(emu_api, name):
    fn, kw = CASES[name]
    got = getattr(S, fn)(emu_api, torch.device("cpu"), **kw)
    compare_outputs(got, load_golden(name), case=name)


@pytest.mark.parametrize("name", ["c1_shear_bgk", "c4_jrl_conc", "c5_jrl_fluid_g"])
def test_case_bit_exact_with_oracle(emu_api, oracle_api, name):
    """Streaming, BGK, moments, moving wall, axisymmetric source: the kernels keep the oracle's
    operation order (csrc/lattice.cuh header), so whole runs agree bit for bit."""
    fn, kw = CASES[name]
    got = getattr(S, fn)(emu_api, torch.device("cpu"), **kw)
    want = getattr(S, fn)(oracle_api, torch.device("cpu"), **kw)
    for k, w in want.items():
        if isinstance(w, np.ndarray):
            assert np.array_equal(np.asarray(got[k]), w, equal_nan=True), k


@pytest.mark.parametrize("name", ["c3_cylinder", "c2_shear_mrt", "ops_plain", "ops_axis_y", "ops_batch2"])
def test_case_close_to_oracle(emu_api, oracle_api, name):
    """MRT (9x9 matvec summation order) and the contact-angle ring differ from the numpy oracle in
    the last bits only."""
    fn, kw = CASES[name]
    got = getattr(S, fn)(emu_api, torch.device("cpu"), **kw)
    want = getattr(S, fn)(oracle_api, torch.device("cpu"), **kw)
    compare_outputs(got, {k: v for k, v in want.items() if v is not None})
