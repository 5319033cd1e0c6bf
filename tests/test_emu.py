"""CPU: the per-read state machines of the CUDA library (dev_*.cuh compiled as C++ by
tests/emu/emu_pipeline.cpp) follow the oracle's control flow exactly -- same SAM bytes."""
import os
import subprocess
import pytest
import common
from common import pkg


@pytest.mark.parametrize("name", ["lambda_default", "syn100_k5", "syn150_sensitive_dpad12", "synN120_default", "syn100_k5_i16", "syn150_sens_i16", "syn100_min300_i16",
                                  "synR100_default", "synR100_k3_vs"])
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


@pytest.mark.parametrize("name", ["lambda_default", "syn100_k5", "syn150_sensitive_dpad12", "synN120_default"])
def test_reference_side_binding_with_oracle_backed_library(name):
    """The drop-in's host half without a GPU: oracle/_ref/bowtie2-align-l-gpu (the reference's driver, FASTQ parser and SAM sink with
    P1-P6 redirected to bt2gpu_align_batch by integration/) run against tests/emu/libbt2gpu_shim.so, which answers the C-ABI with the
    oracle.  Same SAM records as the unmodified binary: the replay through AlnSinkWrap::report and the aliased edit buffer are right.
    (tests/test_gpu_dropin.py runs the same binary against the real library.)"""
    exe = os.path.join(common.REFBIN, "bowtie2-align-l-gpu")
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref/bowtie2-align-l-gpu not built")
    subprocess.run(["make", "-s", "-C", common.ROOT, "emu"], check=True)
    ref, reads, opts, gold = common.CASES[name]
    env = dict(os.environ, BT2GPU="1", BT2GPU_TRACE="1", OMP_NUM_THREADS="4", LD_PRELOAD=os.path.join(common.ROOT, "tests", "emu", "libbt2gpu_shim.so"))
    r = subprocess.run([exe, "-x", common.index_base(ref), "-U", common.reads_path(reads), "-p", "50", "--no-hd"] + opts.split(),
                       env=env, capture_output=True)
    assert r.returncode == 0, r.stderr.decode("latin-1")[-1500:]
    assert b"bt2gpu:" in r.stderr and b"batches" in r.stderr
    assert common.sorted_records(r.stdout.decode("latin-1")) == common.golden_records(gold)
