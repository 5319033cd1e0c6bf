#!/usr/bin/env python
"""FASTQ file -> SAM file through libbt2host.so + libbt2gpu.so on one GPU, stage by stage (host feed and sink).
usage: host_bench.py [n_reads]     (the bench workload: 5 Mbp i.i.d. reference, 150 bp reads, --sensitive --dpad 12)
Prints reads/s of: FASTQ parse + P0 scalars (bt2host_batch_from_fastq), bt2gpu_align_batch (host buffers in, pinned results out),
SAM formatting + write to a file (bt2host_sam_write; also to /dev/null = formatting alone), the three in sequence ("total"), and the three as a pipeline over byte ranges of the file ("pipelined")."""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
    bench = importlib.import_module("bench")
    pkg = importlib.import_module("omp-bowtie2-prime_b200")
    base, codes = bench.make_reference()
    fq = os.path.join(bench.WORK, "host_bench_%d.fq" % n)
    if not os.path.exists(fq):
        bench.write_fastq(fq, bench.simulate_reads(codes, n, seed=2))
    job = pkg.Job(base, bench.OPTS)
    gpu = pkg.GpuAligner(job, 0)
    out = {"reads": n, "fastq_bytes": os.path.getsize(fq), "threads": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count()))}
    best = {}
    for rep in range(3):
        t0 = time.time(); batch = job.fastq(fq); t1 = time.time()
        res = gpu.align(batch); t2 = time.time()
        with open(os.path.join(bench.WORK, "host_bench.sam"), "wb") as f:
            nbytes = job.sam_write(batch, res, f, 1)
        t3 = time.time()
        with open("/dev/null", "wb") as f:      # formatting alone (what the file system adds is the difference to "sam")
            job.sam_write(batch, res, f, 1)
        best["sam_format"] = min(best.get("sam_format", 1e9), time.time() - t3)
        for k, v in (("parse", t1 - t0), ("align", t2 - t1), ("sam", t3 - t2), ("total", t3 - t0)):
            best[k] = min(best.get(k, 1e9), v)
        out["sam_bytes"] = nbytes
        batch.close()
    # the same three stages as a pipeline over byte ranges of the file (omp-bowtie2-prime_b200/stream.py)
    stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    for rep in range(3):            # (one context: two result sets, two block sets)
        t0 = time.time()
        stream.fastq_to_sam(job, fq, os.path.join(bench.WORK, "host_bench.sam"), aligners=[gpu])
        best["pipelined"] = min(best.get("pipelined", 1e9), time.time() - t0)
    for k, v in best.items():
        out[k + "_s"] = round(v, 4); out[k + "_reads_per_s"] = round(n / v)
    print(json.dumps(out))
    gpu.close()


if __name__ == "__main__":
    main()
