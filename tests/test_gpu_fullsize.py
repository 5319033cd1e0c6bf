"""GPU: BASELINE configs[1] at its full size (1 M x 150 bp reads on the 5 Mbp reference, --sensitive --dpad 12) through the
C-ABI, checked with properties that need no oracle run:
  * the results do not depend on how the library cuts the batch (pipeline sub-batch, DP slot count): every struct equal;
  * nearly every simulated read aligns;
  * every alignment's score, gap count, N count and edit count follow from its own edit list (end-to-end scoring: Q40
    mismatch 6, N 1, gap open 5 + 3, extension 3);
  * every gapless alignment, replayed against the reference text at its reported offset, mismatches exactly where its
    edits say, with the reference character they name."""
import os
import sys
import numpy as np
import pytest
import common
from common import pkg

sys.path.insert(0, common.ROOT)

READ_GAP, REF_GAP, MM = 1, 2, 3
N_READS = int(os.environ.get("BT2_TEST_FULL_READS", "1000000"))


def _fields_equal(a, b, skip=("pad", "pad2")):
    assert a.shape == b.shape
    for f in a.dtype.names:
        if f in skip:
            continue
        assert np.array_equal(a[f], b[f]), f


def _workload(n):
    import bench
    base, codes = bench.make_reference()
    reads = bench.simulate_reads(codes, n, seed=2)          # ASCII [n, 150]
    fq = os.path.join(bench.WORK, "fullsize_%d.fq" % n)
    if not os.path.exists(fq):
        bench.write_fastq(fq, reads)
    job = pkg.Job(base, bench.OPTS)
    batch = job.fastq(fq)
    assert batch.n == n
    return job, batch, reads, codes


def check_properties(rd, al, ed, reads, codes, min_aligned):
    n = len(rd)
    # ---- nearly everything aligns (1 % substitutions, one indel in 10 % of the reads)
    assert (rd["n_alns"] > 0).mean() > min_aligned
    na = len(al)
    assert np.array_equal(rd["aln_off"].astype(np.int64), np.cumsum(rd["n_alns"]) - rd["n_alns"])   # alignments are stored in read order

    # ---- score / counts from the alignment's own edits ([edit_off, edit_off + n_edits) of the edit pool)
    ne = al["n_edits"].astype(np.int64)
    owner = np.repeat(np.arange(na), ne)
    idx = np.repeat(al["edit_off"].astype(np.int64), ne) + (np.arange(len(owner)) - np.repeat(np.cumsum(ne) - ne, ne))
    e = ed[idx]
    isn = (e["chr"] == ord("N")) | (e["qchr"] == ord("N"))
    mm = e["type"] == MM
    rdg = e["type"] == READ_GAP
    rfg = e["type"] == REF_GAP
    prev_same = np.zeros(len(e), bool)
    if len(e) > 1:
        same_owner = owner[1:] == owner[:-1]
        prev_same[1:] = same_owner & ((rdg[1:] & rdg[:-1] & (e["pos"][1:] == e["pos"][:-1])) |
                                      (rfg[1:] & rfg[:-1] & (e["pos"][1:] == e["pos"][:-1] + 1)))
    pen = np.where(mm, np.where(isn, 1, 6), np.where(prev_same, 3, 8)).astype(np.float64)
    score = -np.bincount(owner, weights=pen, minlength=na).astype(np.int64)
    assert np.array_equal(score, al["score"].astype(np.int64))
    assert np.array_equal(np.bincount(owner, weights=(rdg | rfg).astype(np.float64), minlength=na).astype(np.int64), al["gaps"].astype(np.int64))
    assert np.array_equal(np.bincount(owner, weights=(mm & isn).astype(np.float64), minlength=na).astype(np.int64), al["ns"].astype(np.int64))
    assert np.array_equal(al["edits"].astype(np.int64), ne)
    assert (al["score"] >= al["minsc"]).all()

    # ---- gapless alignments replayed against the reference text
    read_of = np.repeat(np.arange(n), rd["n_alns"])
    gapless = np.nonzero(al["gaps"] == 0)[0]
    assert len(gapless) > 0.85 * na
    lut = np.full(256, 4, np.uint8); lut[[65, 67, 71, 84]] = [0, 1, 2, 3]
    L = reads.shape[1]
    for lo in range(0, len(gapless), 100000):
        g = gapless[lo:lo + 100000]
        rc = lut[reads[read_of[g]]]
        fw = al["fw"][g] != 0
        rc = np.where(fw[:, None], rc, (3 - rc[:, ::-1]))       # no N in the simulated reads
        off = al["refoff"][g]
        assert (off >= 0).all() and (off + L <= len(codes)).all()
        win = codes[off[:, None] + np.arange(L)[None, :]]
        mism = rc != win
        neg = ne[g]
        assert np.array_equal(mism.sum(axis=1), neg)
        # the edits name the mismatching positions (5'-based for reverse-complement alignments) and the reference characters
        rows, cols = np.nonzero(mism)
        k = np.repeat(al["edit_off"][g].astype(np.int64), neg) + (np.arange(int(neg.sum())) - np.repeat(np.cumsum(neg) - neg, neg))
        eg = ed[k]
        own_row = np.repeat(np.arange(len(g)), neg)
        pos_left = np.where(fw[own_row], eg["pos"].astype(np.int64), L - 1 - eg["pos"].astype(np.int64))
        order = np.lexsort((pos_left, own_row))
        assert np.array_equal(own_row[order], rows) and np.array_equal(pos_left[order], cols)
        assert np.array_equal(eg["chr"][order], np.frombuffer(b"ACGT", np.uint8)[win[rows, cols]])   # forward-strand reference character
        assert np.array_equal(eg["qchr"][order], np.frombuffer(b"ACGT", np.uint8)[rc[rows, cols]])   # character of the aligned strand of the read


def test_property_checker_on_the_oracle():
    """CPU: the checker itself, on results of the oracle for a small prefix of the workload"""
    if not common.have_builder():
        pytest.skip("no builder")
    job, batch, reads, codes = _workload(3000)
    orc = common.Oracle(job)
    r = orc.align(batch, 4)
    check_properties(r.reads, r.alns, r.edits, reads, codes, 0.99)
    orc.close(); batch.close(); job.close()


@pytest.mark.gpu
def test_full_size_properties(monkeypatch):
    job, batch, reads, codes = _workload(N_READS)
    gpu = pkg.GpuAligner(job, 0)
    r = gpu.align(batch)
    rd, al, ed = r.reads.copy(), r.alns.copy(), r.edits.copy()
    gpu.close()
    # ---- independence from the batch cutting
    monkeypatch.setenv("BT2GPU_PIPE", "100003")
    monkeypatch.setenv("BT2GPU_SLOTS", "50000")
    g2 = pkg.GpuAligner(job, 0)
    r2 = g2.align(batch)
    _fields_equal(rd, r2.reads); _fields_equal(al, r2.alns); _fields_equal(ed, r2.edits)
    g2.close()
    check_properties(rd, al, ed, reads, codes, 0.9995)
    batch.close(); job.close()


@pytest.mark.gpu
def test_streamed_fastq_to_sam_equals_one_batch(tmp_path):
    """the reader / aligner / formatter / writer pipeline over byte ranges of the 1 M read file (ONE context: two result sets, two block
    sets) writes the SAM of one batch; then once more with the reader's arrays page-locked (bt2gpu_host_alloc) and the output reserved"""
    import importlib
    stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    job, batch, reads, codes = _workload(N_READS)
    gpu = pkg.GpuAligner(job, 0)
    want = job.sam_header().encode("latin-1") + job.sam_bytes(batch, gpu.align(batch), 1)
    gpu.close()
    import bench
    n, nbytes = stream.fastq_to_sam(job, os.path.join(bench.WORK, "fullsize_%d.fq" % N_READS), str(tmp_path / "s.sam"), chunk_bytes=64 << 20)
    assert n == N_READS
    assert (tmp_path / "s.sam").read_bytes() == want
    pkg.use_pinned_batches()      # (process-wide and once: False if an earlier test did it; arrays made before keep going back to free())
    st = {}
    n, nbytes = stream.fastq_to_sam(job, os.path.join(bench.WORK, "fullsize_%d.fq" % N_READS), str(tmp_path / "p.sam"), chunk_bytes=48 << 20, preallocate=True, stats=st)
    assert n == N_READS and st["ranges"] >= 6
    assert (tmp_path / "p.sam").read_bytes() == want
    batch.close(); job.close()
