"""CPU: the oracle restatement + host sink reproduce, byte for byte, the SAM records the
unmodified reference printed for the committed fixtures (tests/golden/make_golden.sh),
including the reference's aliased-edit CIGAR/MD artefact and YF quirk (F11)."""
import pytest
import common
from common import pkg


@pytest.mark.parametrize("name", sorted(common.CASES))
def test_oracle_matches_reference_golden(name):
    if not common.have_builder():
        pytest.skip("oracle/_ref/bowtie2-build-l not built")
    ref, reads, opts, gold = common.CASES[name]
    job = pkg.Job(common.index_base(ref), opts)
    batch = job.fastq(common.reads_path(reads))
    orc = common.Oracle(job)
    res = orc.align(batch)
    got = common.sorted_records(job.sam(batch, res, mode=1))
    want = common.golden_records(gold)
    assert len(got) == len(want)
    bad = [(g, w) for g, w in zip(got, want) if g != w]
    assert not bad, "first differing record:\n%s\n%s" % bad[0]
    orc.close(); batch.close(); job.close()


def test_oracle_matches_reference_on_the_whole_shipped_example():
    """BASELINE configs[0] in full: all 10 000 reads of example/reads/reads_1.fq on the lambda phage index; the sorted SAM the unmodified
    reference printed here has the md5 the survey recorded (051cc442...) and 6 854 aligned reads"""
    if not common.have_builder():
        pytest.skip("no builder")
    import hashlib, os
    want = common.golden_records("lambda_default_10k.sam")
    assert hashlib.md5(open(os.path.join(common.GOLD, "lambda_default_10k.sam"), "rb").read()).hexdigest() == "051cc442466a220ffbbce47c492a3ca2"
    job = pkg.Job(common.index_base("lambda"), "")
    batch = job.fastq(os.path.join(common.GOLD, "lambda_reads_10k.fq"))
    orc = common.Oracle(job)
    got = common.sorted_records(job.sam(batch, orc.align(batch), mode=1))
    assert got == sorted(want, key=lambda s: s.encode("latin-1"))
    assert sum(1 for l in got if not int(l.split("\t")[1]) & 4) == 6854
    orc.close()


def test_oracle_thread_count_independent():
    if not common.have_builder():
        pytest.skip("no builder")
    job = pkg.Job(common.index_base("syn"), "")
    batch = job.fastq(common.reads_path("syn100"), 0, 300)
    orc = common.Oracle(job)
    a = job.sam(batch, orc.align(batch, 1), 1)
    b = job.sam(batch, orc.align(batch, 4), 1)
    assert a == b


def test_sink_to_file_equals_sink_to_buffer(tmp_path):
    """bt2host_sam_write (blocks written to a descriptor in read order) == bt2host_sam_records, and the summary counters add up"""
    if not common.have_builder():
        pytest.skip("no builder")
    job = pkg.Job(common.index_base("syn"), "-k 3")
    batch = job.fastq(common.reads_path("syn100"), 0, 5000)    # > 2048 reads: more than one block
    orc = common.Oracle(job)
    res = orc.align(batch, 4)
    want = job.sam_bytes(batch, res, 1)
    m0 = job.metrics()
    with open(tmp_path / "o.sam", "wb") as f:
        f.write(b"@HD\n")
        n = job.sam_write(batch, res, f, 1)
    m1 = job.metrics()
    assert n == len(want)
    assert (tmp_path / "o.sam").read_bytes() == b"@HD\n" + want
    assert m1[0] - m0[0] == batch.n
    # a file opened for reading and writing: the blocks go through a shared mapping of the file (twice: growing it, then into a hole)
    with open(tmp_path / "m.sam", "w+b") as f:
        f.write(b"@HD\n")
        assert job.sam_write(batch, res, f, 1) == len(want)
        f.write(b"mid\n")
        assert job.sam_write(batch, res, f, 1) == len(want)
        f.write(b"end\n")
    assert len(want) > (1 << 16)
    assert (tmp_path / "m.sam").read_bytes() == b"@HD\n" + want + b"mid\n" + want + b"end\n"
    # a descriptor that cannot seek (pipe): sequential writes, same bytes
    import os, threading
    r, w = os.pipe()
    got = []
    t = threading.Thread(target=lambda: got.append(os.fdopen(r, "rb").read()))
    t.start()
    with os.fdopen(w, "wb") as f:
        job.sam_write(batch, res, f, 1)
    t.join()
    assert got[0] == want


def test_streamed_fastq_to_sam_equals_whole_file(tmp_path):
    """stream.fastq_to_sam (reader / aligner / sink threads over byte ranges of the file) writes the same SAM as one batch;
    two oracle instances stand in for the two GPU contexts"""
    if not common.have_builder():
        pytest.skip("no builder")
    import importlib
    stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    job = pkg.Job(common.index_base("lambda"), "")
    src = common.reads_path("lambda_reads_1k")
    batch = job.fastq(src)
    o1, o2 = common.Oracle(job), common.Oracle(job)
    want = job.sam_header().encode("latin-1") + job.sam_bytes(batch, o1.align(batch, 4), 1)
    m0 = job.metrics()
    n, nbytes = stream.fastq_to_sam(job, src, str(tmp_path / "s.sam"), chunk_bytes=40000, aligners=[o1, o2])
    got = (tmp_path / "s.sam").read_bytes()
    assert n == 1000 and got == want and nbytes == len(want) - len(job.sam_header().encode("latin-1"))
    assert job.metrics()[0] - m0[0] == 1000
    # with the output file reserved ahead of the sink (and cut back to what was written), explicit reader / sink thread counts
    st = {}
    n, _ = stream.fastq_to_sam(job, src, str(tmp_path / "p.sam"), chunk_bytes=60000, aligners=[o1, o2], preallocate=True, threads=(2, 2), stats=st)
    assert n == 1000 and (tmp_path / "p.sam").read_bytes() == want
    assert st["ranges"] >= 3 and st["threads"] == [2, 2] and st["parse"] > 0 and st["sam"] > 0
    o1.close(); o2.close(); batch.close(); job.close()
