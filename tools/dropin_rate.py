#!/usr/bin/env python3
"""The reference's own binary with its phases P1-P6 on the GPU (oracle/_ref/bowtie2-align-l-gpu, integration/) next to the unmodified
binary, same FASTQ, same options, SAM to a file: wall time, the reference's own phase table, and a byte comparison of the sorted SAM.
usage: python tools/dropin_rate.py [reads]   (needs a GPU and oracle/_ref)"""
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def run(binary, base, fq, out, threads, pmult, gpu):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
    if gpu:
        env["BT2GPU"] = "1"; env["BT2GPU_TRACE"] = "1"
    t0 = time.time()
    r = subprocess.run([os.path.join(bench.REFBIN, binary), "-x", base, "-U", fq, "-S", out, "-p", str(pmult), "-t"] + bench.OPTS.split(), env=env, capture_output=True, text=True)
    dt = time.time() - t0
    if r.returncode != 0:
        raise RuntimeError(binary + " failed: " + r.stderr[-400:])
    phases = {}
    for ln in r.stderr.split("\n"):
        f = ln.split()
        if len(f) == 4 and f[0] == "Timer:" and f[3] != "N/A":
            try:
                phases[f[3]] = round(float(f[2]), 3)
            except ValueError:
                pass
    trace = [ln for ln in r.stderr.split("\n") if ln.startswith("bt2gpu:")]
    return dt, phases, trace


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 400000
    bench.set_config("c2")
    base, codes = bench.make_reference()
    reads = bench.simulate_reads(codes, n, seed=2)
    fq = "/dev/shm/dropin.fq"; bench.write_fastq(fq, reads)
    threads = bench.cpu_cores()
    out = {"reads": n, "threads": threads, "opts": bench.OPTS}
    for tag, binary, gpu, pm in (("reference", "bowtie2-align-l", False, 100), ("dropin_p100", "bowtie2-align-l-gpu", True, 100), ("dropin_p2000", "bowtie2-align-l-gpu", True, 2000)):
        sam = "/dev/shm/dropin_%s.sam" % tag
        dt, phases, trace = run(binary, base, fq, sam, threads, pm, gpu)
        out[tag] = {"wall_s": round(dt, 2), "reads_per_s": round(n / dt), "p": pm, "phase_s": phases, "trace": trace}
    a = sorted(l for l in open("/dev/shm/dropin_reference.sam", "rb") if not l.startswith(b"@"))
    for tag in ("dropin_p100", "dropin_p2000"):
        b = sorted(l for l in open("/dev/shm/dropin_%s.sam" % tag, "rb") if not l.startswith(b"@"))
        out[tag]["records"] = len(b); out[tag]["records_differing"] = sum(1 for x, y in zip(a, b) if x != y) + abs(len(a) - len(b))
    for f in [fq] + ["/dev/shm/dropin_%s.sam" % t for t in ("reference", "dropin_p100", "dropin_p2000")]:
        os.remove(f)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
