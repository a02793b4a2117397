"""CPU (host-emulated kernels): the fused single-launch step equals the solver-object chain."""
import pytest
import torch

from fused_cases import (FUSED_CASES, RUN_CASES, assert_identical, check_run_equals_steps, check_device_tip, check_fluid_walk, check_run_sums_series,
                         check_sums, check_tile_mask, run_pair)


@pytest.mark.parametrize("name", list(FUSED_CASES))
def test_fused_equals_chain(emu_api, name):
    chain, fused = run_pair(emu_api, torch.device("cpu"), name)
    assert_identical(chain, fused)


def test_step_sums(emu_api):
    check_sums(emu_api, torch.device("cpu"))


def test_tile_mask(emu_api):
    check_tile_mask(emu_api, torch.device("cpu"))


@pytest.mark.parametrize("name", list(RUN_CASES))
def test_run_equals_steps(emu_api, oracle_api, name):
    check_run_equals_steps(emu_api, torch.device("cpu"), name, oracle=oracle_api)


def test_run_sums_series(emu_api):
    check_run_sums_series(emu_api, torch.device("cpu"))


@pytest.mark.parametrize("is_negative,fused", [(True, True), (False, True), (True, False)])
def test_device_tip_motion(emu_api, oracle_api, is_negative, fused):
    check_device_tip(emu_api, torch.device("cpu"), is_negative=is_negative, fused=fused, oracle=oracle_api)


@pytest.mark.parametrize("gravity", [2e-5, 0.0])
def test_fluid_walk(emu_api, oracle_api, gravity):
    check_fluid_walk(emu_api, torch.device("cpu"), gravity, oracle=oracle_api)


@pytest.mark.parametrize("geom", ["plane", "45degree"])
def test_device_tip_motion_geometries(emu_api, oracle_api, geom):
    """batch_validation.py's other tips (infinite plane, 45-degree cone) on the device-side tip motion and the walk loop."""
    check_device_tip(emu_api, torch.device("cpu"), is_negative=True, fused=True, oracle=oracle_api, geom=geom, exact=geom != "45degree")


@pytest.mark.parametrize("geom", ["plane", "45degree"])
def test_fluid_walk_geometries(emu_api, oracle_api, geom):
    check_fluid_walk(emu_api, torch.device("cpu"), 2e-5, oracle=oracle_api, geom=geom, exact=geom != "45degree")
