"""Random option sets (seeded): presets, seed length / interval, -D / -R, -k / -M, penalties, minimum score, gap barrier, window padding,
strand switches, trimming, --seed.  CPU: the oracle's SAM == the unmodified reference binary's for every set -- the options are what the
goldens do not vary.  GPU: the CUDA path == the oracle struct for struct on the same sets."""
import os
import random
import subprocess
import pytest
import common
from common import pkg

N_SETS = 96


def option_sets():
    rnd = random.Random(20261105)
    sets = []
    for i in range(N_SETS):
        o = []
        o.append(rnd.choice(["", "--very-fast", "--fast", "--sensitive", "--very-sensitive"]))
        if rnd.random() < 0.4:
            o.append("-L %d" % rnd.choice([10, 14, 18, 20, 22, 25, 28, 32]))
        if rnd.random() < 0.4:
            o.append("-i %s" % rnd.choice(["S,1,0.5", "S,1,1.15", "S,1,2.5", "C,6,0", "C,15,0", "L,1,0.1", "G,1,3"]))
        if rnd.random() < 0.4:
            o.append("-D %d" % rnd.choice([1, 2, 5, 15, 30]))
        if rnd.random() < 0.4:
            o.append("-R %d" % rnd.choice([1, 2, 3, 4]))          # (-R 0 divides by zero in the reference)
        r = rnd.random()
        if r < 0.35:
            o.append("-k %d" % rnd.choice([1, 2, 3, 5, 10]))
        elif r < 0.5:
            o.append("-M %d" % rnd.choice([0, 1, 3, 5]))
        if rnd.random() < 0.35:
            mx = rnd.choice([2, 4, 6, 8]); o.append("--mp %d,%d" % (mx, rnd.choice([1, 2, mx])))
        if rnd.random() < 0.25:
            o.append("--np %d" % rnd.choice([0, 1, 2, 3]))
        if rnd.random() < 0.3:
            o.append("--rdg %d,%d" % (rnd.choice([3, 5, 8]), rnd.choice([1, 2, 3, 4])))
        if rnd.random() < 0.3:
            o.append("--rfg %d,%d" % (rnd.choice([3, 5, 8]), rnd.choice([1, 2, 3, 4])))
        if rnd.random() < 0.4:
            o.append("--score-min %s" % rnd.choice(["L,-0.6,-0.6", "L,0,-0.3", "L,-20,0", "C,-40,0", "L,-0.6,-1.0", "G,-5,-8", "S,-4,-3"]))
        if rnd.random() < 0.25:
            o.append("--gbar %d" % rnd.choice([1, 2, 4, 10, 25]))
        if rnd.random() < 0.5:
            o.append("--dpad %d" % rnd.choice([0, 3, 8, 12, 15]))     # (15 = the default: 150 bp reads then exceed the fork's 200 columns, F13)
        if rnd.random() < 0.15:
            o.append(rnd.choice(["--nofw", "--norc"]))
        if rnd.random() < 0.2:
            o.append("--trim5 %d" % rnd.choice([1, 4, 11]))
        if rnd.random() < 0.2:
            o.append("--trim3 %d" % rnd.choice([2, 7, 30]))
        if rnd.random() < 0.3:
            o.append("--seed %d" % rnd.choice([1, 7, 12345]))
        if rnd.random() < 0.15:
            o.append("--ignore-quals")
        if rnd.random() < 0.2:
            o.append("--n-ceil %s" % rnd.choice(["L,0,0.05", "C,2,0", "L,1,0.5"]))
        reads = rnd.choice(["syn100", "syn150", "synN120", "synR100"])
        if rnd.random() < 0.25:       # the -DENABLE_I16 build on its 16-bit path (--no-sse8), or with a minimum score below the 8-bit range
            o.append("--i16")
            o.append(rnd.choice(["--no-sse8", "--no-sse8", "--score-min L,-280,0"]))
        sets.append((i, reads, " ".join(x for x in o if x)))
    return sets


SETS = option_sets()
REF_OF = {"syn100": "syn", "syn150": "syn", "synN120": "synN", "synR100": "synR"}


def _take(path, n, tmp_path):
    """the first n reads of a FASTQ file (short runs: 24 sets x binary + oracle)"""
    out = str(tmp_path / "sub.fq")
    with open(path) as f, open(out, "w") as g:
        for i, line in enumerate(f):
            if i >= 4 * n:
                break
            g.write(line)
    return out


@pytest.mark.parametrize("idx,reads,opts", SETS, ids=["set%02d" % s[0] for s in SETS])
def test_oracle_equals_reference_binary_on_random_options(idx, reads, opts, tmp_path):
    i16 = "--i16" in opts
    exe = os.path.join(common.REFBIN, "bowtie2-align-l-i16" if i16 else "bowtie2-align-l")
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref/%s not built" % os.path.basename(exe))
    base = common.index_base(REF_OF[reads])
    fq = _take(common.reads_path(reads), 1000, tmp_path)
    r = subprocess.run([exe, "-x", base, "-U", fq, "-p", "30", "--no-hd"] + [x for x in opts.split() if x != "--i16"], capture_output=True,
                       env=dict(os.environ, OMP_NUM_THREADS="2"))
    if r.returncode != 0:
        pytest.skip("the reference rejects this option set: " + r.stderr.decode("latin-1")[-200:])
    want = common.sorted_records(r.stdout.decode("latin-1"))
    job = pkg.Job(base, opts)
    batch = job.fastq(fq)
    orc = common.Oracle(job)
    got = common.sorted_records(job.sam(batch, orc.align(batch), mode=1))
    assert len(got) == len(want) >= 1000
    diff = [(a, b) for a, b in zip(got, want) if a != b]
    assert not diff, "%s\n%d records differ, first:\n%s\n%s" % (opts, len(diff), diff[0][0][:300], diff[0][1][:300])
    orc.close(); batch.close(); job.close()


@pytest.mark.gpu
@pytest.mark.parametrize("idx,reads,opts", SETS, ids=["set%02d" % s[0] for s in SETS])
def test_gpu_equals_oracle_on_random_options(idx, reads, opts, tmp_path):
    base = common.index_base(REF_OF[reads])
    fq = _take(common.reads_path(reads), 1000, tmp_path)
    job = pkg.Job(base, opts)
    batch = job.fastq(fq)
    orc = common.Oracle(job)
    want = orc.align(batch)
    gpu = pkg.GpuAligner(job, 0)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got, i16path="--i16" in opts)
    assert not diffs, opts + "\n" + "\n".join(diffs[:10])
    assert job.sam_bytes(batch, got, 1) == job.sam_bytes(batch, want, 1)
    gpu.close(); orc.close(); batch.close(); job.close()


# ---- the long-read build on reads of 170..255 bp, where the constants-raised reference is a coherent oracle (tests/test_long_reads.py)
LONG_SETS = [(i, o + (" --i16" if "--i16" not in o else "")) for i, _, o in SETS[:32] if "--trim" not in o]


@pytest.mark.parametrize("idx,opts", LONG_SETS, ids=["long%02d" % s[0] for s in LONG_SETS])
def test_long_oracle_equals_constants_raised_reference_on_random_options(idx, opts, tmp_path):
    exe = os.path.join(common.REFBIN, "bowtie2-align-l-long")
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref/bowtie2-align-l-long not built")
    base = common.index_base("syn")
    fq = _take(common.reads_path("synL255"), 1600, tmp_path)
    r = subprocess.run([exe, "-x", base, "-U", fq, "-p", "2", "--no-hd"] + [x for x in opts.split() if x != "--i16"], capture_output=True,
                       env=dict(os.environ, OMP_NUM_THREADS="2"))
    assert r.returncode == 0, r.stderr.decode("latin-1")[-300:]
    want = common.sorted_records(r.stdout.decode("latin-1"))
    job = pkg.Job(base, opts)
    batch = job.fastq(fq)
    orc = common.Oracle(job, long_reads=True)
    got = common.sorted_records(job.sam(batch, orc.align(batch), mode=1))
    diff = [(a, b) for a, b in zip(got, want) if a != b]
    assert len(got) == len(want) and not diff, "%s\n%d records differ, first:\n%s\n%s" % (opts, len(diff), diff[0][0][:300] if diff else "", diff[0][1][:300] if diff else "")
    orc.close(); batch.close(); job.close()


@pytest.mark.gpu
@pytest.mark.parametrize("idx,opts", LONG_SETS, ids=["long%02d" % s[0] for s in LONG_SETS])
def test_long_gpu_equals_oracle_on_random_options(idx, opts, tmp_path):
    base = common.index_base("syn")
    fq = _take(common.reads_path("synLmix") if idx % 2 else common.reads_path("synL255"), 1600, tmp_path)
    job = pkg.Job(base, opts)
    batch = job.fastq(fq)
    orc = common.Oracle(job, long_reads=True)
    want = orc.align(batch)
    gpu = pkg.GpuAligner(job, 0, long_reads=True)
    got = gpu.align(batch)
    diffs = common.compare_results(want, got, i16path=True)
    assert not diffs, opts + "\n" + "\n".join(diffs[:10])
    assert job.sam_bytes(batch, got, 1) == job.sam_bytes(batch, want, 1)
    gpu.close(); orc.close(); batch.close(); job.close()
