"""Reads beyond the reference's 160-row limit (BASELINE configs[4]): the long-read build -- the same sources with
ALN_MAX_ROWS / ALN_MAX_COLS (BT2GPU_MAX_ROWS / BT2GPU_MAX_COLS) raised to 1000 / 1072 -- against oracle D: the golden SAM of
oracle/_ref/bowtie2-align-l-long (the reference with those two constants raised in a scratch copy; NOT the reference as shipped, whose
own answer for these reads is "unaligned") and the restatement compiled with the same constants (oracle/liboracle_long.so).
That binary is a coherent oracle up to 255 bp (cases synL255_*: 170 / 200 / 230 / 255 bp): its seed layer keeps the read length and the
seed offsets in 8-bit fields (aligner_seed.h:679-683, 1226-1230), so a seed at read offset >= 256 is cut from offset mod 256 and framed
on the wrong diagonal.  The restatement and the CUDA path take the offsets as they are; the 250 / 600 / 1000 bp cases (synL_mix_*) still
print the same SAM (the wrapped seeds only add extensions that fail), at scale about 1 % of the reads beyond 255 bp differ."""
import os
import subprocess
import pytest
import common
from common import pkg


@pytest.mark.parametrize("name", sorted(common.CASES_LONG))
def test_long_oracle_matches_constants_raised_reference(name):
    if not common.have_builder():
        pytest.skip("no builder")
    ref, reads, opts, gold = common.CASES_LONG[name]
    job = pkg.Job(common.index_base(ref), opts)
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job, long_reads=True)
    got = common.sorted_records(job.sam(batch, orc.align(batch), mode=1))
    assert got == common.golden_records(gold)
    lens = {len(l.split("\t")[9]) for l in got if not int(l.split("\t")[1]) & 4}
    assert lens >= ({250, 600, 1000} if "mix" in name else {170, 200, 230, 255})   # reads of every length aligned
    orc.close()
    # the build with the reference's own limits leaves every one of them unaligned (SwAligner::initRead fails; F13)
    short = common.Oracle(job)
    res = short.align(batch)
    assert int((res.reads["n_alns"] > 0).sum()) == 0
    short.close()


@pytest.mark.parametrize("name", sorted(common.CASES_LONG))
def test_long_emulated_device_logic_matches_golden(name):
    """dev_*.cuh compiled as C++ with the raised constants (tests/emu/emu_pipeline_long): same SAM bytes"""
    if not common.have_builder():
        pytest.skip("no builder")
    subprocess.run(["make", "-s", "-C", common.ROOT, "emu"], check=True)
    ref, reads, opts, gold = common.CASES_LONG[name]
    out = subprocess.run([os.path.join(common.ROOT, "tests", "emu", "emu_pipeline_long"), common.index_base(ref), common.reads_path(reads), opts, "1"],
                         capture_output=True, check=True)
    assert common.sorted_records(out.stdout.decode("latin-1")) == common.golden_records(gold)


def test_long_fill_lane_code_matches_scalar_restatement():
    subprocess.run(["make", "-s", "-C", common.ROOT, "emu"], check=True)
    out = subprocess.run([os.path.join(common.ROOT, "tests", "emu", "emu_fill_long"), "40", "9"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "lanes == scalar restatement" in out.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(common.CASES_LONG))
def test_long_gpu_matches_oracle_and_golden(name):
    ref, reads, opts, gold = common.CASES_LONG[name]
    job = pkg.Job(common.index_base(ref), opts)
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job, long_reads=True)
    want = orc.align(batch)
    gpu = pkg.GpuAligner(job, 0, long_reads=True)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got, i16path=True)
    assert not diffs, "\n".join(diffs[:10])
    assert common.sorted_records(job.sam(batch, got, mode=1)) == common.golden_records(gold)
    c, oc = gpu.counters(), orc.counters()
    assert c["dp_problems"] == oc["dps"] and c["dp_cells"] == oc["cells"] and c["bt_steps"] == oc["btSteps"]
    gpu.close(); orc.close()


@pytest.mark.gpu
def test_long_and_short_libraries_side_by_side():
    """both builds loaded in one process: each answers for the reads it was built for"""
    job = pkg.Job(common.index_base("syn"), "--i16")
    short_batch = job.fastq(common.reads_path("syn100"), 0, 300)
    long_batch = job.fastq(common.reads_path("synLmix"))
    g_short, g_long = pkg.GpuAligner(job, 0), pkg.GpuAligner(job, 0, long_reads=True)
    a = job.sam_bytes(short_batch, g_short.align(short_batch), 1)
    b = job.sam_bytes(short_batch, g_long.align(short_batch), 1)
    assert a == b                                        # reads within the reference's limits: the same records from either build
    assert int((g_short.align(long_batch).reads["n_alns"] > 0).sum()) == 0
    assert int((g_long.align(long_batch).reads["n_alns"] > 0).sum()) == long_batch.n
    g_short.close(); g_long.close()
