"""GPU: the CUDA path against the UNMODIFIED reference binary run on the same box (oracle/_ref/bowtie2-align-l, -l-i16), sorted SAM records
compared byte for byte, at sizes beyond the committed goldens:
  * a 100 000-read prefix of the bench workload (BASELINE configs[1]: 5 Mbp i.i.d., 150 bp, --sensitive --dpad 12), 8-bit build and the
    -DENABLE_I16 build with --no-sse8;
  * all 10 000 reads of the reference's shipped example (BASELINE configs[0]: lambda phage, reads_1.fq; committed as a fixture).
(tools/scale_check.py does the same at 100 Mbp / 3.1 Gbp; its JSON lines are under profiles/.)"""
import os
import subprocess
import sys
import pytest
import common
from common import pkg

pytestmark = pytest.mark.gpu


def _ref_sam(binary, base, fq, opts, threads=None):
    exe = os.path.join(common.REFBIN, binary)
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref/%s not built" % binary)
    th = threads or min(16, os.cpu_count() or 1)
    env = dict(os.environ, OMP_NUM_THREADS=str(th))
    r = subprocess.run([exe, "-x", base, "-U", fq, "-p", "100", "--no-hd"] + opts.split(), env=env, capture_output=True)
    assert r.returncode == 0, r.stderr.decode("latin-1")[-1000:]
    return sorted(l for l in r.stdout.split(b"\n") if l)


def _gpu_sam(base, fq, opts):
    job = pkg.Job(base, opts)
    batch = job.fastq(fq)
    gpu = pkg.GpuAligner(job, 0)
    res = gpu.align(batch)
    out = sorted(l for l in job.sam_bytes(batch, res, 1).split(b"\n") if l and not l.startswith(b"@"))
    gpu.close()
    return out


def _first_diff(a, b):
    for x, y in zip(a, b):
        if x != y:
            return "\n%s\n%s" % (x.decode("latin-1")[:300], y.decode("latin-1")[:300])
    return "record counts %d vs %d" % (len(a), len(b))


@pytest.fixture(scope="module")
def bench_prefix():
    sys.path.insert(0, common.ROOT)
    import bench
    base, codes = bench.make_reference()
    fq = os.path.join(bench.WORK, "prefix100k.fq")
    bench.write_fastq(fq, bench.simulate_reads(codes, 100000, seed=2))
    return base, fq


def test_bench_workload_prefix_matches_reference_binary(bench_prefix):
    base, fq = bench_prefix
    want = _ref_sam("bowtie2-align-l", base, fq, "--sensitive --dpad 12")
    got = _gpu_sam(base, fq, "--sensitive --dpad 12")
    assert len(want) == 100000
    assert got == want, _first_diff(got, want)


def test_bench_workload_prefix_matches_i16_reference_binary(bench_prefix):
    base, fq = bench_prefix
    want = _ref_sam("bowtie2-align-l-i16", base, fq, "--no-sse8 --sensitive --dpad 12")
    got = _gpu_sam(base, fq, "--i16 --no-sse8 --sensitive --dpad 12")
    assert got == want, _first_diff(got, want)


def test_shipped_example_all_reads_match_reference_binary():
    """BASELINE configs[0]: example/reference/lambda_virus.fa + example/reads/reads_1.fq (10 000 reads, 30-250+ bp, Ns, random qualities):
    6 854 aligned in the reference; against the binary run here and against the committed sorted SAM."""
    fq = os.path.join(common.GOLD, "lambda_reads_10k.fq")
    base = common.index_base("lambda")
    got = _gpu_sam(base, fq, "")
    gold = [l.encode("latin-1") for l in common.golden_records("lambda_default_10k.sam")]
    assert len(gold) == 10000 and sum(1 for l in gold if not int(l.split(b"\t")[1]) & 4) == 6854
    assert got == sorted(gold), _first_diff(got, sorted(gold))
    assert got == _ref_sam("bowtie2-align-l", base, fq, "", threads=4)
