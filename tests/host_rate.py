#!/usr/bin/env python3
"""Host-stage rates on this machine's cores (no GPU): FASTQ range -> SoA batch, and results -> SAM bytes -> file.
The results come from the oracle (test infrastructure; this script is not part of the product path).
usage: python tests/host_rate.py [reads] [workdir]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import common  # noqa: E402
from common import pkg  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    work = sys.argv[2] if len(sys.argv) > 2 else "/dev/shm/bt2_host_rate"
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
