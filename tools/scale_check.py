#!/usr/bin/env python
"""SURVEY 8(d) config C3 at a size one box handles: a large i.i.d. reference (default 100 Mbp: suffix array 400 MB, well
beyond the 126 MB L2), 150 bp simulated reads, `--sensitive --dpad 12`.

  1. builds the .bt2l index with the reference's own builder (oracle/_ref/bowtie2-build-l),
  2. aligns a sample of the reads with the UNMODIFIED reference binary (oracle/_ref/bowtie2-align-l, all host threads) -> SAM,
  3. aligns the same sample through libbt2host.so + libbt2gpu.so -> SAM, and compares the sorted records byte for byte
     (u8 path; with --i16 also the -DENABLE_I16 binary run with --no-sse8 against the library's i16 protocol),
  4. times bt2gpu_align_batch on the full read set (host buffers in, pinned results out).

Prints one JSON line.  Needs a GPU and oracle/_ref (built by oracle/build_ref.sh where the reference sources exist).
--repeats FRAC makes the reference repeat-rich (SURVEY 8d, C4 variant: FRAC of the text covered by copies of 50 families x 300 bp with 2 %
divergence), the workload where -D streaks, several DPs per read, row sampling and MAPQ < 42 occur; --opts replaces the aligner options.
usage: scale_check.py [--ref-mbp 100] [--reads 1000000] [--sample 200000] [--i16] [--repeats 0.1] [--opts "--very-sensitive --dpad 12"] [--tag c4rep]"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REFBIN = os.path.join(ROOT, "oracle", "_ref")
OPTS = "--sensitive --dpad 12"


def sorted_records(sam_bytes):
    return sorted(l for l in sam_bytes.split(b"\n") if l and not l.startswith(b"@"))


def run_ref(binary, base, fq, out, threads, extra, pmult=100):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
    t0 = time.time()
    r = subprocess.run([os.path.join(REFBIN, binary), "-x", base, "-U", fq, "-S", out, "-p", str(pmult)] + OPTS.split() + extra,
                       env=env, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(binary + " failed: " + r.stderr[-300:])
    return time.time() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref-mbp", type=float, default=100.0)
    ap.add_argument("--reads", type=int, default=1000000)
    ap.add_argument("--sample", type=int, default=200000)
    ap.add_argument("--i16", action="store_true")
    ap.add_argument("--work", default=os.path.join(ROOT, "bench_work"))
    ap.add_argument("--repeats", type=float, default=0.0)
    ap.add_argument("--opts", default=None)
    ap.add_argument("--tag", default=None)
    ap.add_argument("--seed", type=int, default=12346)
    ap.add_argument("--keep", action="store_true", help="keep the index and the FASTQ files")
    ap.add_argument("--no-reference", action="store_true", help="skip the reference binary (no parity sample): for runs under ncu")
    ap.add_argument("--long", default=None, help="read lengths, comma separated (e.g. 250,500,750,1000: BASELINE configs[4]): equal shares of each, aligned by "
                    "the long-read build (libbt2gpu_long.so) and, for the parity sample, by oracle/_ref/bowtie2-align-l-long (the -DENABLE_I16 reference "
                    "with ALN_MAX_ROWS / ALN_MAX_COLS raised: not the reference as shipped)")
    args = ap.parse_args()
    global OPTS
    if args.opts:
        OPTS = args.opts
    bench = importlib.import_module("bench")
    pkg = importlib.import_module("omp-bowtie2-prime_b200")
    os.makedirs(args.work, exist_ok=True)
    n_ref = int(args.ref_mbp * 1e6)
    threads = bench.cpu_cores()
    out = {"config": args.tag, "ref_bp": n_ref, "repeat_fraction": args.repeats, "reads": args.reads, "sample": args.sample, "opts": OPTS, "host_threads": threads}

    # ---- reference + index (seed 12346: SURVEY C3)
    base = os.path.join(args.work, "%s%dm" % ("rep" if args.repeats > 0 else "iid", int(args.ref_mbp)))
    rng = np.random.default_rng(args.seed)
    codes = rng.integers(0, 4, size=n_ref, dtype=np.uint8)
    if args.repeats > 0:
        nfam, flen = 50, 300
        fams = rng.integers(0, 4, size=(nfam, flen), dtype=np.uint8)
        ncop = int(args.repeats * n_ref / flen)
        for c0 in range(0, ncop, 65536):                    # copies in blocks: (block, flen) arrays, scattered with one fancy assignment
            nb = min(65536, ncop - c0)
            pos = rng.integers(0, n_ref - flen, size=nb)
            cp = fams[rng.integers(0, nfam, size=nb)].copy()
            m = rng.random((nb, flen)) < 0.02
            cp[m] = (cp[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) % 4
            codes[pos[:, None] + np.arange(flen)[None, :]] = cp
    if not os.path.exists(base + ".4.bt2l"):
        seq = np.frombuffer(b"ACGT", dtype=np.uint8)[codes]
        with open(base + ".fa", "wb") as f:
            f.write(b">seq0\n")
            lines = seq[: (n_ref // 60) * 60].reshape(-1, 60)
            nl = np.full((lines.shape[0], 1), 10, np.uint8)
            f.write(np.concatenate([lines, nl], axis=1).tobytes())
            if n_ref % 60:
                f.write(seq[(n_ref // 60) * 60:].tobytes() + b"\n")
        t0 = time.time()
        subprocess.run([os.path.join(REFBIN, "bowtie2-build-l"), "--threads", str(threads), base + ".fa", base], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out["index_build_s"] = round(time.time() - t0, 1)
    out["index_bytes"] = sum(os.path.getsize(base + e) for e in (".1.bt2l", ".2.bt2l", ".3.bt2l", ".4.bt2l"))

    # ---- reads
    bench.REF_LEN = n_ref
    fq_all = os.path.join(args.work, "scale_all.fq"); fq_s = os.path.join(args.work, "scale_sample.fq")
    long_lens = [int(x) for x in args.long.split(",")] if args.long else None
    ref_binary, job_opts, pmult = "bowtie2-align-l", OPTS, 100
    if long_lens:
        out["read_lengths"] = long_lens
        ref_binary, job_opts, pmult = "bowtie2-align-l-long", OPTS + " --i16", 1   # (-p 1: that build's DP buffers are sized per read slot)
        per, per_s, first = args.reads // len(long_lens), args.sample // len(long_lens), 0
        for k, L in enumerate(long_lens):
            bench.READ_LEN = L
            part = bench.simulate_reads(codes, per, seed=3 + L)
            bench.write_fastq(fq_all, part, first=first, mode="ab" if k else "wb")
            bench.write_fastq(fq_s, part[:per_s], first=first, mode="ab" if k else "wb")
            first += per
        args.reads, args.sample = per * len(long_lens), per_s * len(long_lens)
        out["reads"], out["sample"] = args.reads, args.sample
        out["bases_per_read_mean"] = sum(long_lens) / len(long_lens)
    else:
        reads = bench.simulate_reads(codes, args.reads, seed=3)
        bench.write_fastq(fq_all, reads)
        bench.write_fastq(fq_s, reads[: args.sample])

    # ---- the reference binary on the sample
    sam_ref = os.path.join(args.work, "scale_ref.sam")
    want = None
    if not args.no_reference:
        dt = run_ref(ref_binary, base, fq_s, sam_ref, threads, [], pmult)
        out["reference_binary"] = ref_binary
        out["reference_reads_per_s"] = round(args.sample / dt)
        want = sorted_records(open(sam_ref, "rb").read())

    # ---- this library on the sample
    t0 = time.time()
    job = pkg.Job(base, job_opts)
    out["host_index_load_s"] = round(time.time() - t0, 2)   # .bt2l files -> host arrays (libbt2host; the reference: Ebwt::loadIntoMemory)
    gpu = pkg.GpuAligner(job, 0, long_reads=bool(long_lens))
    out["create_s"] = round(time.time() - t0, 2)      # index load + upload + suffix-array expansion on the device
    out["create_info"] = gpu.create_info()
    batch = job.fastq(fq_s)
    res = gpu.align(batch)
    got = sorted_records(job.sam_bytes(batch, res, 1))
    if want is not None:
        diff = sum(1 for a, b in zip(want, got) if a != b) + abs(len(want) - len(got))
        out["u8_records"] = len(want); out["u8_records_differing"] = diff
        if long_lens:
            # reads beyond 255 bp: the constants-raised reference cuts their seeds from (offset mod 256) -- its seed layer keeps read length and
            # seed offsets in 8-bit fields, aligner_seed.h:679-683 -- and frames them on the wrong diagonal; those records are counted apart
            over = sum(1 for a, b in zip(want, got) if a != b and len(a.split(b"\t")[9]) > 255)
            out["records_differing_reads_over_255bp"] = over
            out["records_differing_reads_up_to_255bp"] = diff - over
            out["note"] = ("parity sample against oracle/_ref/bowtie2-align-l-long (the reference with ALN_MAX_ROWS / ALN_MAX_COLS raised; not the reference as "
                           "shipped); beyond 255 bp that binary's 8-bit seed offsets wrap, differences there are its artefact")
    batch.close()

    # ---- throughput on the full set
    batch = job.fastq(fq_all)
    best = 1e9
    for rep in range(4):
        gpu.counters(reset=True)
        t0 = time.time(); res = gpu.align(batch); dt = time.time() - t0
        if rep:
            best = min(best, dt)
    ctr = gpu.counters()
    out["align_batch_reads_per_s"] = round(args.reads / best)
    out["phase_ms"] = {k: round(ctr[k], 2) for k in ("ms_seed", "ms_resolve", "ms_fill", "ms_bt", "ms_glue")}
    out["sa_probes_per_seed"] = round(ctr["sa_probes"] / max(ctr["seeds_searched"], 1), 2)
    out["dp_problems_per_read"] = round(ctr["dp_problems"] / args.reads, 3)
    out["lf_steps_per_read"] = round((ctr["lf_steps_search"] + ctr["lf_steps_extend"] + ctr["lf_steps_resolve"]) / args.reads, 2)
    out["gcups_fill_kernel"] = round(ctr["dp_cells"] / (ctr["ms_fill"] * 1e-3) / 1e9, 1) if ctr["ms_fill"] > 0 else None
    if args.reads <= (1 << 20):   # inputs resident in HBM: device time of the whole hot path (CUDA events)
        gpu.upload(batch); gpu.run_resident(); gpu.counters(reset=True)
        for _ in range(3):
            gpu.run_resident()
        c2 = gpu.counters()
        out["resident_reads_per_s"] = round(3 * args.reads / (c2["ms_run"] * 1e-3))
        out["resident_phase_ms_per_run"] = {k: round(c2[k] / 3, 2) for k in ("ms_seed", "ms_resolve", "ms_fill", "ms_bt", "ms_glue", "ms_run")}
    out["aligned_fraction"] = float((res.reads["n_alns"] > 0).sum()) / args.reads
    batch.close(); gpu.close()

    # ---- the 16-bit protocol
    if args.i16 and os.access(os.path.join(REFBIN, "bowtie2-align-l-i16"), os.X_OK):
        run_ref("bowtie2-align-l-i16", base, fq_s, sam_ref, threads, ["--no-sse8"])
        want = sorted_records(open(sam_ref, "rb").read())
        job16 = pkg.Job(base, OPTS + " --i16 --no-sse8")
        gpu16 = pkg.GpuAligner(job16, 0)
        batch = job16.fastq(fq_s)
        got = sorted_records(job16.sam_bytes(batch, gpu16.align(batch), 1))
        out["i16_records"] = len(want)
        out["i16_records_differing"] = sum(1 for a, b in zip(want, got) if a != b) + abs(len(want) - len(got))
        batch.close(); gpu16.close()
    if not args.keep:
        for f in (fq_all, fq_s, sam_ref):
            if os.path.exists(f):
                os.remove(f)
    print(json.dumps(out))
    bad = out.get("records_differing_reads_up_to_255bp", 0) if long_lens else out.get("u8_records_differing", 0)
    return 0 if bad == 0 and out.get("i16_records_differing", 0) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
