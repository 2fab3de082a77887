"""GPU: the reference's two 100-year time loops with the device behind them (scripts/gpu_time_loop.py holds the device backend of the loop model).

(1) Halfar dome, SIA + semi-implicit thickness steps, and (2) MISMIP_mod, DIVA + thickness steps + masks + sub-grid fractions: every call the loop makes
goes through the narrow entries of the C ABI; per predictor-corrector iteration the run must equal the oracle-backed loop (tests/test_reference_halfar_log.py,
tests/test_reference_mismip_mod_log.py, both pinned on the reference's checksum logs) and follow the logs themselves.
(3) MISMIP_mod again with ONE ufe_b200_pc_step per predictor-corrector step (geometry stage on the device, the reference's loop semantics): it must land on
the loop model's end state to 1e-6 -- the step the bench times is the reference's step.
"""
import importlib.util
import os

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def tl():
    spec = importlib.util.spec_from_file_location("gpu_time_loop", os.path.join(ROOT, "scripts", "gpu_time_loop.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def mismip_cpu(tl):
    return tl.mismip_mod_cpu_loop()


def test_halfar_time_loop_through_the_narrow_seam(tl):
    assert tl.halfar_narrow_seam()


def test_mismip_mod_time_loop_through_the_narrow_seam(tl, mismip_cpu):
    assert tl.mismip_mod_narrow_seam(*mismip_cpu)


def test_mismip_mod_one_call_per_step_lands_on_the_reference_loop(tl, mismip_cpu):
    assert tl.one_call_mismip_mod(mismip_cpu[0])
