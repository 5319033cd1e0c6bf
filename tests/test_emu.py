"""CPU: the per-read state machines of the CUDA library (dev_*.cuh compiled as C++ by
tests/emu/emu_pipeline.cpp) follow the oracle's control flow exactly -- same SAM bytes."""
import os
import subprocess
import pytest
import common
from common import pkg


@pytest.mark.parametrize("name", ["lambda_default", "syn100_k5", "syn150_sensitive_dpad12", "synN120_default", "syn100_k5_i16", "syn150_sens_i16"])
def test_emulated_device_logic_matches_golden(name):
    if not common.have_builder():
        pytest.skip("no builder")
    exe = os.path.join(common.ROOT, "tests", "emu", "emu_pipeline")
    subprocess.run(["make", "-s", "-C", common.ROOT, "emu"], check=True)
    ref, reads, opts, gold = common.CASES[name]
    out = subprocess.run([exe, common.index_base(ref), common.reads_path(reads), opts, "1"], capture_output=True, check=True)
    got = common.sorted_records(out.stdout.decode("latin-1"))
    assert got == common.golden_records(gold)


def test_fill_kernel_lane_code_matches_scalar_restatement():
    """K_fill's lane code (k_fill.cuh through the C++ stand-ins of simd16.cuh), 32 lanes stepped in lock step, for K = 2, 4, 8, 16,
    banded and whole-rectangle: flags, H/E/F, last row, candidates == the scalar restatement the emulated pipeline uses."""
    subprocess.run(["make", "-s", "-C", common.ROOT, "emu"], check=True)
    out = subprocess.run([os.path.join(common.ROOT, "tests", "emu", "emu_fill"), "120", "5"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "lanes == scalar restatement" in out.stdout
