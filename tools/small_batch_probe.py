#!/usr/bin/env python3
"""bt2gpu_align_batch on small batches (what a host loop with few read slots hands over): wall time per call and the phase table.
usage: python tools/small_batch_probe.py"""
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    bench.set_config("c2")
    pkg = importlib.import_module("omp-bowtie2-prime_b200")
    base, codes = bench.make_reference()
    job = pkg.Job(base, bench.OPTS)
    reads = bench.simulate_reads(codes, 65536, seed=2)
    fq = "/dev/shm/small.fq"; bench.write_fastq(fq, reads)
    gpu = pkg.GpuAligner(job, 0)
    for n in (256, 1600, 8192, 65536):
        b = job.fastq(fq, 0, n)
        for _ in range(5):
            gpu.align(b)
        gpu.counters(reset=True)
        reps = 40
        t0 = time.perf_counter()
        for _ in range(reps):
            gpu.align(b)
        dt = (time.perf_counter() - t0) / reps
        c = gpu.counters()
        print("%6d reads: %.3f ms per call (%.2f M reads/s); per call: run %.3f (seed %.3f glue %.3f step %.3f fill %.3f bt %.3f) h2d %.3f d2h %.3f ms; %d launches" % (
            n, dt * 1e3, n / dt / 1e6, c["ms_run"] / reps, c["ms_seed"] / reps, c["ms_glue"] / reps, c["ms_resolve"] / reps, c["ms_fill"] / reps, c["ms_bt"] / reps,
            c["ms_h2d"] / reps, c["ms_d2h"] / reps, c["launches"] // reps), flush=True)
        b.close()
    gpu.close(); job.close(); os.remove(fq)


if __name__ == "__main__":
    main()
