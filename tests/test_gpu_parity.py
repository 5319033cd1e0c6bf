"""GPU: the CUDA path, called through the C-ABI (libbt2gpu.so), equals the oracle
struct-for-struct (alignments, edits, RNG state, counters, F14 scratch) and reproduces
the reference's golden SAM byte for byte."""
import numpy as np
import pytest
import common
from common import pkg

pytestmark = pytest.mark.gpu


def _gpu(job):
    return pkg.GpuAligner(job, 0)


@pytest.mark.parametrize("name", sorted(common.CASES))
def test_gpu_matches_oracle_and_golden(name):
    ref, reads, opts, gold = common.CASES[name]
    job = pkg.Job(common.index_base(ref), opts)
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got, i16path="--i16" in opts and ("--no-sse8" in opts or "--score-min" in opts))
    assert not diffs, "\n".join(diffs[:10])
    sam = common.sorted_records(job.sam(batch, got, mode=1))
    assert sam == common.golden_records(gold)
    c = gpu.counters()
    oc = orc.counters()
    assert c["dp_problems"] == oc["dps"] and c["dp_cells"] == oc["cells"]
    # seeds are searched through the device's full suffix array (binary search) where it exists; the rest take the reference's LF walk
    assert c["lf_steps_search"] <= oc["lfSearch"] and (c["sa_probes"] > 0 or c["lf_steps_search"] == oc["lfSearch"])
    # single-row extensions are text comparisons through the full suffix array, one per reference LF step
    assert c["lf_steps_extend"] + c["ext_text_steps"] == oc["lfExtend"]
    assert c["bt_steps"] == oc["btSteps"] and c["sa_rows_resolved"] == oc["saRows"]
    assert c["launches"] > 0
    gpu.close(); orc.close()


@pytest.mark.parametrize("name", ["syn100_default", "syn100_k5", "lambda_default"])
def test_gpu_whole_rectangle_fill(name):
    """--exact-xs (bt2gpu_params.exactBestLtMinsc = 1): K_fill computes whole rectangles like the reference's EEU8_alignNucleotides and
    prm.bestLtMinscMate1 -- the one field the banded fill only bounds from below (checked as such in the other tests) -- is equal too;
    everything else is what the banded fill gives (same SAM bytes), and the kernel reports having computed more cells."""
    ref, reads, opts, gold = common.CASES[name]
    job = pkg.Job(common.index_base(ref), (opts + " --exact-xs").strip())
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got, exact_xs=True)
    assert not diffs, "\n".join(diffs[:10])
    assert common.sorted_records(job.sam(batch, got, mode=1)) == common.golden_records(gold)
    cx = gpu.counters()
    gpu.close()
    jb = pkg.Job(common.index_base(ref), opts)
    gb = _gpu(jb)
    gb.align(batch)
    cb = gb.counters()
    assert cb["dp_cells"] == cx["dp_cells"] and cb["dp_cells_computed"] < cx["dp_cells_computed"]
    gb.close(); orc.close()


@pytest.mark.parametrize("name", ["lambda_k3_vs", "syn150_sensitive_dpad12"])
def test_gpu_without_full_suffix_array(name, monkeypatch):
    """BT2GPU_FULL_SA=0: the path an index of 2^32 rows or more takes -- every seed through the reference's LF walk (k_seed_walk),
    offsets through LF walks to the sampled suffix array.  Same results, and exactly the oracle's LF step counts."""
    monkeypatch.setenv("BT2GPU_FULL_SA", "0")
    ref, reads, opts, gold = common.CASES[name]
    job = pkg.Job(common.index_base(ref), opts)
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got)
    assert not diffs, "\n".join(diffs[:10])
    assert common.sorted_records(job.sam(batch, got, mode=1)) == common.golden_records(gold)
    c = gpu.counters(); oc = orc.counters()
    assert c["sa_probes"] == 0 and c["ext_text_steps"] == 0
    assert c["lf_steps_search"] == oc["lfSearch"] and c["lf_steps_extend"] == oc["lfExtend"]
    assert c["lf_steps_resolve"] == oc["lfResolve"] if "lfResolve" in oc else True
    gpu.close(); orc.close()


def test_gpu_edge_cases():
    """empty read, read shorter than the seed, all-N read, read > 160 bp (dropped like the reference), Ns inside seeds"""
    job = pkg.Job(common.index_base("syn"), "")
    fa = open(common.index_base("syn") + ".fa").read().split("\n")
    s = "".join(l for l in fa[1:200] if not l.startswith(">"))
    seqs = ["", s[100:110], "N" * 50, s[1000:1170], s[2000:2100], s[3000:3050] + "N" + s[3051:3100],
            s[4000:4022], s[5000:5023], "ACGT" * 30, s[6000:6160]]
    quals = ["I" * len(x) for x in seqs]
    batch = job.arrays(seqs, quals)
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got)
    assert not diffs, "\n".join(diffs[:10])
    assert job.sam(batch, got, 1) == job.sam(batch, want, 1)
    gpu.close(); orc.close()


@pytest.mark.parametrize("ktab", ["", "12"])
def test_gpu_reads_at_the_end_of_the_text(ktab, monkeypatch):
    """seeds that match the last bases of the joined text: the suffixes there run off the end, which the index sorts AFTER
    every base (the suffix-array search has to agree); with the device prefix table forced on for a small index as well
    (BT2GPU_KTAB=12) and its self-check on (BT2GPU_CHECK_INDEX)"""
    if ktab:
        monkeypatch.setenv("BT2GPU_KTAB", ktab)
    monkeypatch.setenv("BT2GPU_CHECK_INDEX", "1")
    comp = {"A": "T", "C": "G", "G": "C", "T": "A", "N": "N"}
    for ref in ("syn", "lambda"):
        job = pkg.Job(common.index_base(ref), "")
        fa = open(common.index_base(ref) + ".fa").read() if ref != "lambda" else open(common.os.path.join(common.GOLD, "lambda_virus.fa")).read()
        last = "".join(fa.split(">")[-1].split("\n")[1:]).upper()     # the last sequence: its end is the end of the joined text
        seqs = []
        for end in range(0, 12):                                       # reads ending 0..11 bases before the end, three lengths
            for L in (60, 100, 150):
                r = last[len(last) - end - L: len(last) - end]
                seqs += [r, "".join(comp[c] for c in reversed(r))]
        seqs += [last[-22:], last[-30:], last[-23:-1]]
        batch = job.arrays(seqs, ["I" * len(x) for x in seqs])
        orc = common.Oracle(job)
        want = orc.align(batch)
        gpu = _gpu(job)
        got = gpu.align(batch)
        diffs = common.compare_results(want, got)
        assert not diffs, ref + ": " + "\n".join(diffs[:10])
        assert job.sam(batch, got, 1) == job.sam(batch, want, 1)
        assert (got.reads["n_alns"] > 0).mean() > 0.9
        gpu.close(); orc.close()


def test_gpu_batch_split_and_determinism(monkeypatch):
    """results do not depend on the internal batch size or the number of DP slots (SURVEY 8b determinism contract)"""
    job = pkg.Job(common.index_base("syn"), "-k 2")
    batch = job.fastq(common.reads_path("syn100"), 0, 700)
    g1 = _gpu(job)
    a = g1.align(batch)
    sam_a = job.sam(batch, a, 1)
    g1.close()
    monkeypatch.setenv("BT2GPU_BATCH", "256")
    monkeypatch.setenv("BT2GPU_SLOTS", "100")
    g2 = _gpu(job)
    b = g2.align(batch)
    assert job.sam(batch, b, 1) == sam_a
    # split API == align_batch
    monkeypatch.setenv("BT2GPU_BATCH", "4096")
    g3 = _gpu(job)
    g3.upload(batch); g3.run_resident(); c = g3.fetch()
    assert job.sam(batch, c, 1) == sam_a
    g2.close(); g3.close()


def test_gpu_pool_overflow_grows_and_reruns(monkeypatch):
    """-k 20 on the repeat-rich reference with the device pools started 400x too small (BT2GPU_POOL_DIV): the batch overflows them, the
    library doubles the pool that ran out and runs the batch again until it fits -- same results as the oracle, no "pool exhausted" error."""
    monkeypatch.setenv("BT2GPU_POOL_DIV", "400")
    job = pkg.Job(common.index_base("syn"), "-k 20 -D 5 -R 3")
    batch = job.fastq(common.reads_path("syn100"))
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got)
    assert not diffs, "\n".join(diffs[:10])
    assert job.sam(batch, got, 1) == job.sam(batch, want, 1)
    assert int(got.reads["n_alns"].max()) > 6          # more alignments per read than the default pools budget
    gpu.close(); orc.close()


def test_gpu_larger_iid_vs_oracle():
    """5 Mbp i.i.d. reference, 150 bp reads, --sensitive --dpad 12 (BASELINE config 2 at reduced read count)"""
    base = common.index_base("iid5000000")
    import os
    fq = os.path.join(common.WORK, "iid150.fq")
    if not os.path.exists(fq):
        common.gen(["reads", base + ".fa", fq, 20000, 150, "--seed", 2])
    job = pkg.Job(base, "--sensitive --dpad 12")
    batch = job.fastq(fq)
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = _gpu(job)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got)
    assert not diffs, "\n".join(diffs[:10])
    assert job.sam(batch, got, 1) == job.sam(batch, want, 1)
    gpu.close(); orc.close()


def test_gpu_i16_score_range_limit():
    """-DENABLE_I16 protocol: a minimum score below -254 selects the reference's 16-bit path (golden syn100_min300_i16: alignments down to
    AS:i:-299).  The fill keeps scores down to -28 416; a minimum score beyond that fails loudly instead of returning clamped results."""
    job = pkg.Job(common.index_base("syn"), "--i16 --score-min L,-30000,0")
    batch = job.fastq(common.reads_path("syn100"), 0, 50)
    gpu = _gpu(job)
    with pytest.raises(RuntimeError, match="16-bit"):
        gpu.align(batch)
    gpu.close()


def test_submit_wait_and_result_lifetime():
    """bt2gpu_submit / bt2gpu_wait == bt2gpu_align_batch, and what a call returned is still there after the next call on the same
    context (two sets of result arrays taken in turn): one context is enough for a reader / aligner / sink pipeline"""
    job = pkg.Job(common.index_base("syn"), "-k 3")
    a = job.fastq(common.reads_path("syn100"), 0, 700)
    b = job.fastq(common.reads_path("syn150"), 0, 900)
    gpu = _gpu(job)
    ra = gpu.align(a)
    sam_a = job.sam_bytes(a, ra, 1)
    gpu.submit(b)
    with pytest.raises(RuntimeError, match="in flight"):
        gpu.align(a)                       # one batch in flight per context
    assert job.sam_bytes(a, ra, 1) == sam_a  # read while b is being aligned
    rb = gpu.wait()
    sam_b = job.sam_bytes(b, rb, 1)
    assert job.sam_bytes(a, ra, 1) == sam_a  # ... and after
    g2 = _gpu(job)
    assert job.sam_bytes(b, g2.align(b), 1) == sam_b
    with pytest.raises(RuntimeError, match="nothing was submitted"):
        gpu.wait()
    gpu.submit(a); rc = gpu.wait()          # takes the set ra lived in
    assert job.sam_bytes(a, rc, 1) == sam_a and job.sam_bytes(b, rb, 1) == sam_b
    gpu.close(); g2.close()
