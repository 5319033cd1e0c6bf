#!/usr/bin/env python3
"""Where the FASTQ -> SAM pipeline spends its time on this box: the three stages one after the other on the same ranges (nothing
overlapping), then the pipeline itself, with the host library's per-call trace.  usage: python tools/f2f_probe.py [reads] [repeat]"""
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
    repeat = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    bench.set_config("c2")
    pkg = importlib.import_module("omp-bowtie2-prime_b200"); stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    base, codes = bench.make_reference()
    job = pkg.Job(base, bench.OPTS)
    reads = bench.simulate_reads(codes, n, seed=2)
    fq, sam = "/dev/shm/probe.fq", "/dev/shm/probe.sam"
    for r in range(repeat):
        bench.write_fastq(fq, reads, first=r * n, mode="ab" if r else "wb")
    size = os.path.getsize(fq)
    gpu = pkg.GpuAligner(job, 0)
    chunk = 64 << 20
    for threads in ((16, 16), (5, 9)):
        job.set_threads(*threads)
        for rep in range(2):
            tp = ta = ts = 0.0; off = 0; nr = 0
            with open(sam, "w+b") as f:
                while off < size:
                    t0 = time.perf_counter(); b, nxt = job.fastq_range(fq, off, chunk); t1 = time.perf_counter()
                    if b.n == 0:
                        b.close(); break
                    res = gpu.align(b); t2 = time.perf_counter()
                    job.sam_write(b, res, f, 1); t3 = time.perf_counter()
                    nr += b.n; b.close(); off = nxt
                    tp += t1 - t0; ta += t2 - t1; ts += t3 - t2
            print("serial, threads %s: %d reads  parse %.3f s (%.1f M/s)  align %.3f s (%.1f M/s)  sam %.3f s (%.1f M/s)" % (
                threads, nr, tp, nr / tp / 1e6, ta, nr / ta / 1e6, ts, nr / ts / 1e6), flush=True)
    # one range, looked at closely: the sink right after the aligner returned, then again on the same results
    job.set_threads(0, 0)
    b, _ = job.fastq_range(fq, 0, chunk)
    for rep in range(3):
        res = gpu.align(b)
        with open("/dev/null", "wb") as f:
            ts = []
            for k in range(3):
                t0 = time.perf_counter(); job.sam_write(b, res, f, 1); ts.append((time.perf_counter() - t0) * 1e3)
        time.sleep(0.2)
        with open("/dev/null", "wb") as f:
            t0 = time.perf_counter(); job.sam_write(b, res, f, 1); t_idle = (time.perf_counter() - t0) * 1e3
        print("one range (%d reads): sink right after align %.1f ms, again %.1f ms, again %.1f ms, after 0.2 s idle %.1f ms" % (b.n, ts[0], ts[1], ts[2], t_idle), flush=True)
    b.close()
    for out in (sam, "/dev/null"):
        for rep in range(2):
            st = {}
            t0 = time.perf_counter(); nr, nb = stream.fastq_to_sam(job, fq, out, 0, aligners=[gpu], stats=st); dt = time.perf_counter() - t0
            print("pipeline -> %s: %.3f s  %.2f M reads/s  %s" % (out, dt, nr / dt / 1e6, st), flush=True)
    gpu.close(); job.close()
    for f in (fq, sam):
        os.remove(f)


if __name__ == "__main__":
    main()
