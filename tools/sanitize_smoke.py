#!/usr/bin/env python
"""A smoke-sized batch through the C-ABI, for `compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_smoke.py [nreads]`
(the reference's analogue is its -fsanitize build, Makefile:192-199).  Checks the results against the oracle as well,
so that a sanitizer run is also a parity run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pkg = common.pkg
for name, reads, opts in (("lambda", "lambda_reads_1k", ""), ("syn", "syn100", "-k 5 --very-sensitive")):
    job = pkg.Job(common.index_base(name), opts)
    batch = job.fastq(common.reads_path(reads), 0, n)
    gpu = pkg.GpuAligner(job, 0)
    got = gpu.align(batch)
    orc = common.Oracle(job)
    diffs = common.compare_results(orc.align(batch), got)
    print("%s: %d reads, %d differences vs the oracle" % (name, batch.n, len(diffs)))
    gpu.close(); orc.close()
    if diffs:
        sys.exit(1)
