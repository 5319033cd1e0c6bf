#!/usr/bin/env python
"""Debug helper (GPU box): run one case through the CUDA library and the oracle and print every difference.
usage: gpu_check.py <ref-name> <reads-name|path.fq> "<opts>" [max_reads]"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import common
from common import pkg

ref, reads, opts = sys.argv[1], sys.argv[2], sys.argv[3]
maxn = int(sys.argv[4]) if len(sys.argv) > 4 else 0
job = pkg.Job(common.index_base(ref), opts)
batch = job.fastq(reads if reads.endswith(".fq") else common.reads_path(reads), 0, maxn)
orc = common.Oracle(job)
want = orc.align(batch)
gpu = pkg.GpuAligner(job, 0)
got = gpu.align(batch)
d = common.compare_results(want, got)
print("\n".join(d) if d else "IDENTICAL", flush=True)
print("oracle", orc.counters())
print("gpu   ", dict(gpu.counters()))
for f in ("nonzTot", "numElts", "numRanges", "nSeedsTried", "nExIters", "nRedundants"):
    import numpy as np
    bad = np.nonzero(want.reads[f] != got.reads[f])[0]
    print(f, len(bad), [(int(i), int(want.reads[f][i]), int(got.reads[f][i])) for i in bad[:5]])
import numpy as np
names = want.reads.dtype.names
bad = [i for i in range(len(want.reads)) if any(want.reads[f][i] != got.reads[f][i] for f in names if f not in ("aln_off", "scratch_off"))]
print("differing reads:", len(bad), bad[:20])
r = batch.reads.contents
import ctypes as C
off = np.ctypeslib.as_array(C.cast(r.seq_off, C.POINTER(C.c_uint32)), shape=(r.n + 1,))
for i in bad[:4]:
    print("---- read", i, "len", off[i + 1] - off[i])
    for nm, R in (("oracle", want), ("gpu", got)):
        rr = R.reads[i]
        print(nm, {f: int(rr[f]) for f in names})
        for k in range(int(rr["n_alns"])):
            a = R.alns[rr["aln_off"] + k]
            print("   aln", {f: int(a[f]) for f in a.dtype.names})
