#!/usr/bin/env python3
"""Host-stage rates on this machine's cores (no GPU): FASTQ range -> SoA batch, and results -> SAM bytes -> file.
The results come from the oracle (test infrastructure; this script is not part of the product path).
With --gpu the results come from the CUDA path instead, and the stages are also timed at several thread counts.
usage: python tests/host_rate.py [reads] [workdir] [--gpu]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import common  # noqa: E402
from common import pkg  # noqa: E402


def main():
    gpu = "--gpu" in sys.argv
    argv = [a for a in sys.argv if a != "--gpu"]
    n = int(argv[1]) if len(argv) > 1 else 200000
    work = argv[2] if len(argv) > 2 else "/dev/shm/bt2_host_rate"
    os.makedirs(work, exist_ok=True)
    fa, fq, base = os.path.join(work, "r.fa"), os.path.join(work, "r_%d.fq" % n), os.path.join(work, "r")
    if not os.path.exists(base + ".1.bt2l"):
        common.gen(["ref", fa, 2000000, "--seed", 5])
        common._run([os.path.join(common.REFBIN, "bowtie2-build-l"), "--threads", "8", fa, base])
    if not os.path.exists(fq):
        common.gen(["reads", fa, fq, n, 150, "--seed", 6])
    job = pkg.Job(base, "--sensitive")
    size = os.path.getsize(fq)
    for rep in range(3):
        t0 = time.time(); b, _ = job.fastq_range(fq, 0, size); t1 = time.time()
        print("parse: %d reads %.1f ms  %.2f M reads/s  %.2f GB/s" % (b.n, (t1 - t0) * 1e3, b.n / (t1 - t0) / 1e6, size / (t1 - t0) / 1e9))
        if rep < 2:
            b.close()
    if gpu:
        import ctypes
        omp = ctypes.CDLL("libgomp.so.1")
        orc = pkg.GpuAligner(job, 0)
        res = orc.align(b)
        out = os.path.join(work, "out.sam")
        for nt in (1, 2, 4, 8, 16):
            omp.omp_set_num_threads(nt)
            t0 = time.time(); b2, _ = job.fastq_range(fq, 0, size); t1 = time.time(); b2.close()
            t2 = time.time(); txt = job.sam_bytes(b, res, 1); t3 = time.time()
            with open(out, "wb") as f:
                t4 = time.time(); nb = job.sam_write(b, res, f, 1); t5 = time.time()
            with open(out, "w+b") as f:
                t8 = time.time(); nb = job.sam_write(b, res, f, 1); t9 = time.time()
            with open(out, "r+b") as f:          # pages of the file exist already
                t6 = time.time(); nb = job.sam_write(b, res, f, 1); t7 = time.time()
            print("threads %2d: parse %.1f ms (%.2f M/s)  sam format+copy %.1f ms (%.2f M/s)  format+write new file %.1f ms (%.2f M/s)  rewrite %.1f ms  new file, mapped %.1f ms (%.2f M/s)" % (
                nt, (t1 - t0) * 1e3, b.n / (t1 - t0) / 1e6, (t3 - t2) * 1e3, b.n / (t3 - t2) / 1e6, (t5 - t4) * 1e3, b.n / (t5 - t4) / 1e6, (t7 - t6) * 1e3, (t9 - t8) * 1e3, b.n / (t9 - t8) / 1e6))
        # the same results copied out of the pinned arrays into ordinary memory: does the sink read pinned memory slower?
        import numpy as np
        raw = res.raw
        keep = []
        r2 = pkg.bt2gpu_results(); r2.n = raw.n; r2.n_alns = raw.n_alns; r2.n_edits = raw.n_edits
        for name, cnt, typ in (("reads", raw.n, pkg.bt2gpu_read_result), ("alns", raw.n_alns, pkg.bt2gpu_aln), ("edits", raw.n_edits, pkg.bt2gpu_edit)):
            nbytes = max(1, int(cnt)) * ctypes.sizeof(typ)
            a = np.empty(nbytes, dtype=np.uint8); ctypes.memmove(a.ctypes.data, ctypes.cast(getattr(raw, name), ctypes.c_void_p), int(cnt) * ctypes.sizeof(typ))
            keep.append(a); setattr(r2, name, ctypes.cast(a.ctypes.data, ctypes.POINTER(typ)))
        res2 = pkg.Results(r2)
        for nt in (1, 4, 16):
            omp.omp_set_num_threads(nt)
            for which, rr in (("pinned", res), ("pageable", res2)):
                with open("/dev/null", "wb") as f:
                    t0 = time.time(); job.sam_write(b, rr, f, 1); t1 = time.time()
                print("threads %2d, results in %s memory -> /dev/null: %.1f ms (%.2f M/s)" % (nt, which, (t1 - t0) * 1e3, b.n / (t1 - t0) / 1e6), flush=True)
        orc.close(); b.close(); job.close()
        return
    orc = common.Oracle(job)
    t0 = time.time(); res = orc.align(b); print("oracle align %.1f s" % (time.time() - t0))
    out = os.path.join(work, "out.sam")
    for rep in range(3):
        with open(out, "wb") as f:
            t0 = time.time(); nb = job.sam_write(b, res, f, 1); t1 = time.time()
        print("sam: %d bytes %.1f ms  %.2f M reads/s  %.2f GB/s" % (nb, (t1 - t0) * 1e3, b.n / (t1 - t0) / 1e6, nb / (t1 - t0) / 1e9))
    for rep in range(3):
        t0 = time.time(); txt = job.sam_bytes(b, res, 1); t1 = time.time()
        print("sam format only: %d bytes %.1f ms  %.2f M reads/s" % (len(txt), (t1 - t0) * 1e3, b.n / (t1 - t0) / 1e6))
    orc.close(); b.close(); job.close()


if __name__ == "__main__":
    main()
