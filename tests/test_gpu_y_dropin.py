"""GPU: the drop-in built for real.  oracle/_ref/bowtie2-align-l-gpu is the REFERENCE's own driver, FASTQ parser and
SAM sink (bt2_search.cpp, pat.cpp, aln_sink.cpp, sam.cpp compiled from /root/reference by oracle/build_ref.sh) with phases P1-P6 of its
batch loop replaced by bt2gpu_align_batch through integration/gpu_batch.cpp + integration/bt2_search_gpu.patch.  Its SAM must be the
unmodified binary's (the golden files), record for record: the reference's aln_sink.cpp reproduces the bytes when fed bt2gpu_results,
including the CIGAR / MD:Z it reads through the aliased per-slot edit buffer."""
import os
import subprocess
import pytest
import common

pytestmark = pytest.mark.gpu
BIN = os.path.join(common.REFBIN, "bowtie2-align-l-gpu")


def _run(name, threads, pmult, env_extra=None):
    ref, reads, opts, gold = common.CASES[name]
    env = dict(os.environ, BT2GPU="1", OMP_NUM_THREADS=str(threads))
    env.update(env_extra or {})
    r = subprocess.run([BIN, "-x", common.index_base(ref), "-U", common.reads_path(reads), "-p", str(pmult), "--no-hd"] + opts.split(),
                       env=env, capture_output=True, text=True, encoding="latin-1")
    assert r.returncode == 0, r.stderr[-2000:]
    return common.sorted_records(r.stdout), r.stderr


@pytest.mark.parametrize("name", ["lambda_default", "lambda_k3_vs", "syn150_sensitive_dpad12", "syn100_k5", "synN120_default"])
def test_patched_reference_binary_matches_unmodified_one(name):
    if not os.access(BIN, os.X_OK):
        pytest.skip("oracle/_ref/bowtie2-align-l-gpu not built (oracle/build_ref.sh where /root/reference exists)")
    got, err = _run(name, 4, 64, {"BT2GPU_TRACE": "1"})
    assert "bt2gpu:" in err and "batches" in err, "the GPU path did not run:\n" + err[-500:]
    assert got == common.golden_records(common.CASES[name][3])


def test_patched_reference_binary_is_batch_size_independent():
    """F10: the multiset of records does not depend on threads x -p (here: the size of the batches handed to the GPU)"""
    if not os.access(BIN, os.X_OK):
        pytest.skip("no patched binary")
    a, _ = _run("syn100_k5", 2, 7)
    b, _ = _run("syn100_k5", 8, 200)
    assert a == b == common.golden_records(common.CASES["syn100_k5"][3])
