#!/usr/bin/env python
"""Headline benchmark of the seed-and-extend hot path (BASELINE.json: aligned reads/sec,
150 bp, --sensitive end-to-end; DP GCUPS beside it).

  python bench.py --gpus N --steps K --warmup W          CUDA path (libbt2gpu.so through the C-ABI)
  python bench.py --impl reference ...                   the reference's own OpenMP CPU path (oracle/_ref)

Workload = BASELINE configs[1]: synthetic 5 Mbp i.i.d. reference, 1 M simulated 150 bp reads
(1 % substitutions, one 1-bp indel in 10 % of reads), unpaired, `--sensitive --dpad 12`
(the reference's 200-column DP buffer cannot hold a 150 bp read at the default --dpad 15,
SURVEY F13).  One step = one pass of the whole hot path over one batch of reads; every
rank aligns its own batch against its own replica of the index (weak scaling, no collective
on the data path).  `value` times bt2gpu_run_resident (inputs resident in HBM) with CUDA
events recorded on the library's stream; `e2e` times bt2gpu_align_batch from host buffers
(H2D + kernels + D2H of the results).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
WORK = os.path.join(ROOT, "bench_work")
REFBIN = os.path.join(ROOT, "oracle", "_ref")
OPTS = "--sensitive --dpad 12"
REF_LEN = 5_000_000
# --config: the workloads of SURVEY 8(d).  c2 (default) is BASELINE configs[1], the one the metric is quoted on; the others are the larger
# references as one GPU takes them (index built on first use with the reference's builder: 30 s at 100 Mbp, 16 min at 3.1 Gbp)
CONFIGS = {
    "c2":    {"name": "configs[1]", "ref": "5 Mbp i.i.d. reference", "len": 5_000_000, "seed": 12345, "repeats": 0.0, "opts": "--sensitive --dpad 12", "base": "iid5m"},
    "c3":    {"name": "configs[2] (unpaired, one GPU)", "ref": "100 Mbp i.i.d. reference", "len": 100_000_000, "seed": 12346, "repeats": 0.0, "opts": "--sensitive --dpad 12", "base": "iid100m"},
    "c4":    {"name": "configs[3] (one GPU)", "ref": "3.1 Gbp i.i.d. reference", "len": 3_100_000_000, "seed": 12347, "repeats": 0.0, "opts": "--very-sensitive --dpad 12", "base": "iid3100m"},
    "c4rep": {"name": "configs[3], repeat-rich variant at 100 Mbp", "ref": "100 Mbp reference, 10 % of it 50 repeat families x 300 bp", "len": 100_000_000, "seed": 12346, "repeats": 0.1,
              "opts": "--very-sensitive --dpad 12", "base": "rep100m"},
}
CFG = CONFIGS["c2"]


def set_config(name):
    global CFG, OPTS, REF_LEN
    CFG = CONFIGS[name]; OPTS = CFG["opts"]; REF_LEN = CFG["len"]
READ_LEN = 150
METRIC = "aligned_reads_per_sec_150bp_sensitive_e2e"
ALU_WARP_INST_PER_CLK_SM = 1.95   # VIMNMX.S16x2 / HSET2 / LOP3 share this pipe (measured, profiles/r1_pipe_bench.txt)
# ALU-pipe instructions a pair of DP cells needs in K_fill (one warp-instruction = 32 lanes x 2 problems = 64 cells): 3 VIMNMX(3).S16x2 of the
# recurrence + 1 HSET2 (mismatch mask) + 1 LOP3 (mask & penalty) + 7 HSET2 (the backtrace's comparisons) + 1 LOP3 (merge of the carried flags)
ALU_INST_PER_CELL_PAIR = 13
# profiles/r2_ncu_fill_raw.csv: dram__bytes_read.sum + dram__bytes_write.sum of one K_fill launch over 262 144 DP problems, per problem;
# profiles/r1_ncu_full_raw.csv for k_bt
NCU_TRAFFIC_PER_PROBLEM = {"k_fill": (0.368e9 + 5.85e9) / 262144, "k_bt": 5.68e9 / 262144}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def make_reference():
    """The config's synthetic reference (seeded) + its .bt2l index, built once with the reference's own builder."""
    os.makedirs(WORK, exist_ok=True)
    base = os.path.join(WORK, CFG["base"])
    fa = base + ".fa"
    rng = np.random.default_rng(CFG["seed"])
    codes = rng.integers(0, 4, size=REF_LEN, dtype=np.uint8)
    if CFG["repeats"] > 0:      # copies of 50 families x 300 bp with 2 % divergence until that fraction of the text is covered
        nfam, flen = 50, 300
        fams = rng.integers(0, 4, size=(nfam, flen), dtype=np.uint8)
        ncop = int(CFG["repeats"] * REF_LEN / flen)
        for c0 in range(0, ncop, 65536):
            nb = min(65536, ncop - c0)
            pos = rng.integers(0, REF_LEN - flen, size=nb)
            cp = fams[rng.integers(0, nfam, size=nb)].copy()
            m = rng.random((nb, flen)) < 0.02
            cp[m] = (cp[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) % 4
            codes[pos[:, None] + np.arange(flen)[None, :]] = cp
    if not os.path.exists(base + ".4.bt2l"):
        seq = np.frombuffer(b"ACGT", dtype=np.uint8)[codes]
        with open(fa, "wb") as f:
            f.write(b">seq0\n")
            lines = seq[: (REF_LEN // 60) * 60].reshape(-1, 60)
            f.write(np.concatenate([lines, np.full((lines.shape[0], 1), 10, np.uint8)], axis=1).tobytes())
            if REF_LEN % 60:
                f.write(seq[(REF_LEN // 60) * 60:].tobytes() + b"\n")
        builder = os.path.join(REFBIN, "bowtie2-build-l")
        if not os.access(builder, os.X_OK):
            raise RuntimeError("oracle/_ref/bowtie2-build-l missing: run oracle/build_ref.sh where /root/reference exists")
        subprocess.run([builder, "--threads", str(min(16, os.cpu_count() or 1)), fa, base], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return base, codes


def simulate_reads(codes, n, seed):
    """n reads of READ_LEN bp: uniform start, 1 % substitutions, a 1-bp indel in 10 % of the reads, random strand, Q40."""
    if n > 100000:  # bound the temporaries
        parts = [simulate_reads(codes, min(100000, n - i), seed * 7919 + i) for i in range(0, n, 100000)]
        return np.concatenate(parts, axis=0)
    rng = np.random.default_rng(seed)
    L = READ_LEN
    st = rng.integers(0, REF_LEN - L - 2, size=n)
    win = codes[st[:, None] + np.arange(L + 2)[None, :]]
    sub = rng.random((n, L + 2)) < 0.01
    win = np.where(sub, (win + rng.integers(1, 4, size=(n, L + 2), dtype=np.uint8)) % 4, win).astype(np.uint8)
    indel = np.nonzero(rng.random(n) < 0.10)[0]
    pos = rng.integers(10, L - 10, size=len(indel))
    ins = rng.random(len(indel)) < 0.5
    newb = rng.integers(0, 4, size=len(indel), dtype=np.uint8)
    for k, r in enumerate(indel):
        p = pos[k]
        if ins[k]:
            win[r, p + 1:] = win[r, p:-1].copy(); win[r, p] = newb[k]
        else:
            win[r, p:-1] = win[r, p + 1:].copy()
    rd = win[:, :L]
    rc = rng.random(n) < 0.5
    rd = np.where(rc[:, None], 3 - rd[:, ::-1], rd).astype(np.uint8)
    return np.frombuffer(b"ACGT", dtype=np.uint8)[rd]  # ASCII [n, L]


def write_fastq(path, ascii_reads, first=0, mode="wb"):
    n, L = ascii_reads.shape
    q = b"I" * L
    with open(path, mode) as f:
        for i in range(n):
            f.write(b"@r%d\n" % (first + i) + ascii_reads[i].tobytes() + b"\n+\n" + q + b"\n")
        return f.tell()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_ = gpu, [], threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_.is_set():
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self.stop_.wait(0.2)

    def summary(self):
        self.stop_.set()
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        reasons = []
        for i, nm in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
            if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows):
                reasons.append(nm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def run_reference_binary(base, fq, nreads, threads, pmult):
    """One run of the unmodified reference aligner on a FASTQ sample.  Returns wall time (index load, FASTQ parse, all phases and
    SAM formatting to /dev/null included) and the reference's own MyTimer phase table (bt2_search.cpp:2244-2280, printed to
    stderr; it is reset at batch iteration 4, so the first 3 iterations -- 3 x threads x p reads -- are not in it)."""
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
    t0 = time.time()
    r = subprocess.run([os.path.join(REFBIN, "bowtie2-align-l"), "-x", base, "-U", fq, "-S", "/dev/null", "-p", str(pmult), "-t"] + OPTS.split(),
                       env=env, capture_output=True, text=True)
    dt = time.time() - t0
    if r.returncode != 0:
        raise RuntimeError("reference aligner failed: " + r.stderr[-300:])
    phases = {}
    search_s = None
    for ln in r.stderr.split("\n"):
        f = ln.split()
        if len(f) == 4 and f[0] == "Timer:" and f[3] != "N/A":
            try:
                phases[f[3]] = float(f[2])
            except ValueError:
                pass
        elif ln.startswith("Multiseed full-index search:"):   # -t, whole seconds
            h, m, sec = ln.split(":", 1)[1].strip().split(":")
            search_s = int(h) * 3600 + int(m) * 60 + int(sec)
    hot = sum(phases.get(k, 0.0) for k in HOT_PHASES)
    timed_reads = max(nreads - 3 * threads * pmult, 0)      # reads the phase table covers (one seed round per read on this workload)
    return {"wall_s": dt, "search_s": search_s, "phases": phases, "hot_path_s": hot, "extendSeeds_s": phases.get("extendSeeds"),
            "timed_reads": timed_reads, "threads": threads, "p": pmult, "reads": nreads}


# the phases bt2gpu_align_batch replaces (P1-P6, bt2_search.cpp:2591-2805); "read" (P0) and "finishReadOne" (P7) stay on the host
HOT_PHASES = ("instantiateSeeds", "searchAllSeedsPrepare", "searchAllSeedsDo", "searchAllSeedsFinalize", "prioritizeSATups", "extendSeeds")
CELLS_PER_READ = 150 * 199   # one DP per read on this workload, rdlen x rflen


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def tune_reference(base, codes, cores, log_=True, quick=False):
    """-p is a batch multiplier, not a thread count, and must be tuned: sweep -p in {30,100,300,1000} (and 16 threads
    vs all cores when the box has more) on a small sample and keep the configuration with the best hot-path rate."""
    n = int(os.environ.get("BT2_BENCH_TUNE_READS", "60000"))
    fq = os.path.join(WORK, "ref_tune.fq")
    write_fastq(fq, simulate_reads(codes, n, seed=3))
    tlist = sorted({cores, min(cores, 16)})
    if os.environ.get("BT2_BENCH_REF_P"):
        plist = [int(os.environ["BT2_BENCH_REF_P"])]
    else:
        plist = [30, 100, 300, 1000]
    if quick:                       # the cpu_baseline leg of the CUDA arm: a bounded sweep (the full one is the reference arm's)
        tlist, plist = [min(cores, 16)], [p for p in plist if p in (100, 300)] or plist[:1]
    sweep, best = [], None
    for th in tlist:
        for pm in plist:
            if 4 * th * pm > n:       # the phase table would cover less than a quarter of the sample
                continue
            r = run_reference_binary(base, fq, n, th, pm)
            rate_hot = r["timed_reads"] / r["hot_path_s"] if r["hot_path_s"] > 0 else 0.0
            rate_wall = n / r["wall_s"]
            sweep.append({"threads": th, "p": pm, "wall_reads_per_s": round(rate_wall), "hot_path_reads_per_s": round(rate_hot)})
            if log_:
                log("[bench] reference sweep: threads %d -p %d: wall %.0f reads/s, hot path (P1-P6) %.0f reads/s" % (th, pm, rate_wall, rate_hot))
            if best is None or rate_hot > best[0]:
                best = (rate_hot, th, pm)
    if best is None:
        best = (0.0, cores, 100)
    return best[1], best[2], sweep


def summarize_reference(runs, sweep, cores):
    """cpu_baseline block from timed runs of one configuration."""
    wall = sum(r["wall_s"] for r in runs); reads = sum(r["reads"] for r in runs)
    hot = sum(r["hot_path_s"] for r in runs); treads = sum(r["timed_reads"] for r in runs)
    ext = sum(r["extendSeeds_s"] or 0.0 for r in runs)
    ph = {}
    for r in runs:
        for k, v in r["phases"].items():
            ph[k] = ph.get(k, 0.0) + v
    out = {"unit": "reads/s", "cores": cores, "kind": "reference", "threads": runs[0]["threads"], "p": runs[0]["p"],
           "wall_s": wall, "wall_reads_per_s": reads / wall if wall > 0 else None,
           "search_s": sum(r["search_s"] or 0 for r in runs),
           "hot_path_s": hot, "hot_path_reads_per_s": treads / hot if hot > 0 else None,
           "extendSeeds_s": ext, "cpu_gcups": treads * CELLS_PER_READ / ext / 1e9 if ext > 0 and CFG["opts"] == "--sensitive --dpad 12" else None,   # (one DP per read there)
           "phase_s": {k: round(v, 4) for k, v in ph.items()}, "phase_table_reads": treads, "p_sweep": sweep}
    return out


def reference_arm(args, rank, world):
    """--impl reference: the reference's own OpenMP CPU implementation on the host cores (rank 0 only).
    `value` is the like-for-like of the CUDA arm's `e2e`: reads / (sum of the reference's own phase timers P1-P6), i.e. the hot path
    alone with the reads already parsed into host memory and no SAM formatting -- the most favourable reading for the CPU.  The
    whole-process figure (start-up, FASTQ parse, SAM formatting to /dev/null) is cpu_baseline.wall_reads_per_s."""
    if rank != 0:
        return
    base, codes = make_reference()
    cores = cpu_cores()
    threads, pmult, sweep = tune_reference(base, codes, cores)
    # per-step sample of the 1 M-read workload, sized so that the whole run stays within a few minutes
    budget_s = float(os.environ.get("BT2_BENCH_REF_BUDGET_S", "240"))
    rate = max([x["wall_reads_per_s"] for x in sweep if x["threads"] == threads and x["p"] == pmult] + [10000])
    sample = int(os.environ.get("BT2_BENCH_REF_READS", "0")) or int(min(1_000_000, max(100_000, budget_s * rate / max(args.steps + args.warmup, 1))))
    sample = max(sample, 8 * threads * pmult)
    reads = simulate_reads(codes, sample, seed=2)
    fq = os.path.join(WORK, "ref_sample.fq")
    write_fastq(fq, reads)
    for _ in range(args.warmup):
        run_reference_binary(base, fq, sample, threads, pmult)
    runs = [run_reference_binary(base, fq, sample, threads, pmult) for _ in range(args.steps)]
    cb = summarize_reference(runs, sweep, cores)
    val = cb["hot_path_reads_per_s"]
    cb["value"] = val
    cb["sample"] = ("%d of the workload's 1M reads per step x %d steps, OMP_NUM_THREADS=%d -p %d (best of the sweep), OMP_PROC_BIND=close; value = reads covered by "
                    "the reference's phase timers / sum of its P1-P6 phase times" % (sample, args.steps, threads, pmult))
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * sample / val if val else None,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": workload_config(),
            "cpu_baseline": cb,
            "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(n=None):
    """The same `config` object for both arms (the reference arm runs a bounded per-step sample of it, stated in cpu_baseline.sample)."""
    return {"workload": "%s: %s, 1M x 150 bp simulated reads (1 %% substitutions, 1-bp indel in 10 %%), unpaired, %s" % (CFG["name"], CFG["ref"], OPTS),
            "reads_per_step_per_gpu": int(n if n is not None else os.environ.get("BT2_BENCH_READS", "1000000")), "parallelism": "reads sharded, index replicated per GPU",
            "l2": "inputs larger than L2 (300 MB of reads per step, DP cell flags written per problem; 126 MB L2)"}


def file_to_file(pkg, job, ascii_reads, device, repeat=4):
    """FASTQ file -> SAM file on one GPU (omp-bowtie2-prime_b200/stream.py: reader thread, aligner, sink thread), both files on
    tmpfs: the like-for-like of the reference process's wall time (cpu_baseline.wall_reads_per_s: FASTQ file -> SAM to /dev/null).
    The file holds the workload's reads `repeat` times over (own names), so that the three-stage pipeline runs in its steady state."""
    stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") else WORK
    fq, sam = os.path.join(tmp, "bt2gpu_bench.fq"), os.path.join(tmp, "bt2gpu_bench.sam")
    n = ascii_reads.shape[0]
    for r in range(repeat):
        write_fastq(fq, ascii_reads, first=r * n, mode="ab" if r else "wb")
    try:
        al = [pkg.GpuAligner(job, device)]
        stream.fastq_to_sam(job, fq, sam, device, aligners=al)          # warm-up (page cache, pinned buffers)
        best, stages = None, None
        for _ in range(2):
            st = {}
            t0 = time.time()
            nr, nb = stream.fastq_to_sam(job, fq, sam, device, aligners=al, stats=st)
            dt = time.time() - t0
            if best is None or dt < best:
                best, stages = dt, st
        # the same with the SAM text thrown away, as the reference arm's wall time is taken (-S /dev/null): what the pipeline does when
        # the page allocation of a 1.6 GB tmpfs file (the kernel's work, one page at a time under the file's lock) is out of the way
        null_best, null_stages = None, None
        for _ in range(2):
            st = {}
            t0 = time.time()
            stream.fastq_to_sam(job, fq, "/dev/null", device, aligners=al, stats=st)
            dt = time.time() - t0
            if null_best is None or dt < null_best:
                null_best, null_stages = dt, st
        for g in al:
            g.close()
        busy = {k: round(stages[k], 4) for k in ("parse", "align", "sam", "write")}
        null_busy = {k: round(null_stages[k], 4) for k in ("parse", "align", "sam", "write")}
        return {"value": nr / best, "unit": "reads/s", "reads": nr, "fastq_bytes": os.path.getsize(fq), "sam_bytes": nb, "seconds": best,
                "where": tmp, "host_cores": cpu_cores(), "ranges": stages["ranges"], "reader_sink_threads": stages["threads"], "stage_busy_s": busy, "limiting_stage": max(busy, key=busy.get),
                "sam_to_dev_null": {"value": nr / null_best, "unit": "reads/s", "seconds": null_best, "stage_busy_s": null_busy, "limiting_stage": max(null_busy, key=null_busy.get)},
                "note": "FASTQ file -> parse -> bt2gpu_align_batch -> SAM formatting -> SAM file, pipelined over byte ranges of the file; stage_busy_s = "
                        "seconds each of the four overlapping stages was at work (parse, sam = formatting and write share the host cores)"}
    finally:
        for f in (fq, sam):
            if os.path.exists(f):
                os.remove(f)


def file_to_file_sharded(pkg, job, ascii_reads, device, rank, world, dist, torch):
    """SURVEY 8(e) as a file-to-file run on N GPUs of one box: ONE FASTQ file on tmpfs (every rank appends its reads in rank order), the
    byte range of shard r goes to rank r (its own reader / aligner / sink pipeline, index replicated), every rank writes its part of the
    SAM, parts are concatenated in shard order.  No collective on the data path; time = barrier to barrier, max over ranks.  All ranks
    share the host's cores for parsing and formatting, which is what limits it (file_to_file at N = 1 is already host-bound)."""
    stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") else WORK
    fq = os.path.join(tmp, "bt2gpu_bench_all.fq")
    n = ascii_reads.shape[0]
    ends = torch.zeros(world, dtype=torch.int64, device="cuda")
    for r in range(world):                       # ranks append in turn: one file, shard boundaries = what each rank wrote
        if r == rank:
            ends[r] = write_fastq(fq, ascii_reads, first=r * n, mode="wb" if r == 0 else "ab")
        dist.barrier()
    dist.all_reduce(ends, op=dist.ReduceOp.SUM)
    ends = [int(x) for x in ends.tolist()]
    start, end = (ends[rank - 1] if rank else 0), ends[rank]
    part = os.path.join(tmp, "bt2gpu_bench.sam.%d" % rank)
    share = max(2, cpu_cores() // world)           # the ranks of one box split its cores
    thr = (max(1, share * 3 // 8), max(1, share * 3 // 8))
    al = [pkg.GpuAligner(job, device)]
    try:
        stream.fastq_to_sam(job, fq, part, device, aligners=al, start=start, end=end, header=rank == 0, threads=thr)     # warm-up
        best = None
        for _ in range(2):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.time()
            nr, nb = stream.fastq_to_sam(job, fq, part, device, aligners=al, start=start, end=end, header=rank == 0, threads=thr)
            dist.barrier()
            dt = torch.tensor([time.time() - t0], dtype=torch.float64, device="cuda")
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            best = float(dt.item()) if best is None else min(best, float(dt.item()))
        tot = torch.tensor([nr, nb], dtype=torch.int64, device="cuda")
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        return {"value": int(tot[0].item()) / best, "unit": "reads/s", "reads": int(tot[0].item()), "fastq_bytes": ends[-1], "sam_bytes": int(tot[1].item()),
                "seconds": best, "where": tmp, "shards": world, "host_cores": cpu_cores(), "reader_sink_threads_per_rank": list(thr),
                "note": "one FASTQ file, one byte range per rank -> parse -> bt2gpu_align_batch -> SAM formatting -> one SAM part per rank (concatenated in shard "
                        "order); limited by the host cores all ranks share for parsing and formatting"}
    finally:
        for g in al:
            g.close()
        dist.barrier()
        for f in ([fq] if rank == 0 else []) + [part]:
            if os.path.exists(f):
                os.remove(f)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="bt2gpu")
    ap.add_argument("--reads", type=int, default=int(os.environ.get("BT2_BENCH_READS", "1000000")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-file-to-file", action="store_true")
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    set_config(args.config)
    if args.impl == "reference":
        return reference_arm(args, rank, world)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the bt2gpu path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("omp-bowtie2-prime_b200")
    if os.environ.get("BT2_BENCH_PINNED", "1") != "0":
        pkg.use_pinned_batches()      # the host library's batch arrays are page-locked: the file-to-file pipeline's aligner call copies them straight up

    if rank == 0:
        base, codes = make_reference()
    if world > 1:
        dist.barrier()
    if rank != 0:
        base, codes = make_reference()
    job = pkg.Job(base, OPTS)
    n = args.reads
    t0 = time.time()
    ascii_reads = simulate_reads(codes, n, seed=2 + 1000 * rank)
    off = (np.arange(n + 1, dtype=np.uint64) * READ_LEN).astype(np.uint32)
    seqb = ascii_reads.tobytes(); qualb = b"I" * (n * READ_LEN)
    err = C.create_string_buffer(512)
    bp = job._h.bt2host_batch_from_arrays(job.ptr, n, off.ctypes.data, seqb, qualb, None, err, 512)
    if not bp:
        raise RuntimeError(err.value.decode())
    batch = pkg.Batch(job, bp)
    log("[bench] rank %d: %d reads simulated and packed in %.1fs" % (rank, n, time.time() - t0))
    gpu = pkg.GpuAligner(job, local)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident timing: inputs already in HBM
    gpu.upload(batch)
    for _ in range(args.warmup):
        gpu.run_resident()
    gpu.counters(reset=True)
    sampler = ClockSampler(local); sampler.start()
    sync_all()
    w0 = time.time()
    for _ in range(args.steps):
        gpu.run_resident()
    sync_all()
    wall = time.time() - w0
    ctr = gpu.counters(reset=True)
    clocks = sampler.summary()
    dev_s = ctr["ms_run"] / 1000.0  # CUDA events on the library's stream around every run_resident
    res = gpu.fetch()
    aligned = int((res.reads["n_alns"] > 0).sum())
    # ---- end to end through the C-ABI from (pinned) host buffers: H2D + all kernels + D2H of the results
    src = batch.reads.contents
    pinned = []
    preads = pkg.bt2gpu_reads(); preads.n = n
    for name, nbytes in (("seq_off", 4 * (n + 1)), ("seq", n * READ_LEN), ("qual", n * READ_LEN), ("rng_seed", 4 * n),
                         ("minsc", 4 * n), ("nceil", 4 * n), ("interval", 4 * n)):
        t = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        C.memmove(t.data_ptr(), getattr(src, name), nbytes)
        pinned.append(t)
        setattr(preads, name, t.data_ptr())
    for _ in range(args.warmup):           # (both of the context's result sets grow to the batch's size)
        gpu.align_raw(preads)
    gpu.counters(reset=True)
    sync_all()
    e0 = time.time()
    for _ in range(args.steps):
        r2 = gpu.align_raw(preads)
    sync_all()
    e2e_s = time.time() - e0
    d2h = n * C.sizeof(pkg.bt2gpu_read_result) + int(r2.raw.n_alns) * C.sizeof(pkg.bt2gpu_aln) + int(r2.raw.n_edits) * C.sizeof(pkg.bt2gpu_edit)
    h2d = 2 * n * READ_LEN + 4 * (n + 1) + 16 * n

    t_dev = torch.tensor([dev_s, e2e_s, wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    dev_s, e2e_s, wall = [float(x) for x in t_dev.tolist()]
    total_reads = n * world * args.steps
    value = total_reads / dev_s
    peak, peak_src = peaks()
    # dominant kernel by device time, algorithmic bytes per launch (DESIGN.md "Kernels"):
    #   k_seed  8 B per prefix-table lookup + 16 B per suffix-array probe (4 B gather + 12 B of text) + 128 B per distinct BWT side touched by an LF step
    #           (multi-row left extension) + 4 B SA gather and 1 B of text per compared base of a single-row extension
    #   k_step  4 B SA gather per resolved row + ~60 B of packed reference and 208 B of window written per DP problem
    #   k_fill  1 B of flags per DP cell (stored as 8-byte units of 5 cells for 129..160 bp reads) + ~0.9 KB of inputs/outputs per problem
    #   k_bt    1 B of flags read + 1 B written (visited mark) per backtrace step
    phases = {"k_seed": (ctr["ms_seed"], ctr["n_seed"], 128.0 * ctr["lf_sides_touched"] + 8.0 * ctr["seeds_searched"] + 16.0 * ctr["sa_probes"] + 1.25 * ctr["ext_text_steps"]),
              "k_step": (ctr["ms_resolve"], ctr["n_step"], 128.0 * ctr["lf_steps_resolve"] + 4.0 * ctr["sa_rows_resolved"] + 268.0 * ctr["dp_problems"]),
              "k_fill": (ctr["ms_fill"], ctr["n_fill"], 1.0 * ctr["dp_cells"] + 900.0 * ctr["dp_problems"]),
              "k_bt": (ctr["ms_bt"], ctr["n_bt"], 2.0 * ctr["bt_steps"])}
    dom = max(phases, key=lambda k: phases[k][0])
    # K_fill and k_bt are within a few per cent of each other on the headline workload: the kernel reported is K_fill (the one with a
    # ceiling to be measured against) unless another kernel is clearly ahead; k_bt's own figures go to roofline["k_bt"] either way
    if phases["k_fill"][0] >= 0.9 * phases[dom][0]:
        dom = "k_fill"
    ms, nl, byts = phases[dom]
    if dom == "k_fill":   # what the fill stores: one flag byte per cell it COMPUTES (the band) + its inputs / outputs
        byts = 1.0 * ctr["dp_cells_computed"] + 900.0 * ctr["dp_problems"]
    hbm_achieved = (byts / max(nl, 1)) / ((ms / max(nl, 1)) * 1e-3) / 1e9 if ms > 0 else 0.0
    # ncu --set full capture of the same command (profiles/): dram bytes read+written by the dominant kernel, scaled from the captured
    # launch to this run's average launch (problems per launch), so that it compares with the algorithmic bytes per launch
    traffic = NCU_TRAFFIC_PER_PROBLEM[dom] * ctr["dp_problems"] / max(nl, 1) if dom in NCU_TRAFFIC_PER_PROBLEM else None
    # K_fill against the integer ALU pipe.  GCUPS counts the cells the REFERENCE computes for the same DP problems
    # (rdlen x rflen each); the kernel computes only the band of diagonals an alignment >= minsc can reach (cells_computed)
    alu_peak = 148 * 1.965e9 * ALU_WARP_INST_PER_CLK_SM * 64.0 / ALU_INST_PER_CELL_PAIR / 1e9
    fill_gcups = ctr["dp_cells"] / (ctr["ms_fill"] * 1e-3) / 1e9 if ctr["ms_fill"] > 0 else 0.0
    fill_gcups_computed = ctr["dp_cells_computed"] / (ctr["ms_fill"] * 1e-3) / 1e9 if ctr["ms_fill"] > 0 else 0.0
    hbm = {"achieved": hbm_achieved, "peak": peak, "unit": "GB/s", "frac": hbm_achieved / peak, "peak_source": peak_src, "algorithmic_bytes_per_launch": byts / max(nl, 1)}
    if dom == "k_fill":
        roofline = {"bound": "int_alu", "kernel": dom, "achieved": fill_gcups, "peak": alu_peak, "unit": "GCUPS", "frac": fill_gcups / alu_peak,
                    "traffic": traffic, "launches": int(nl), "ms_per_launch": ms / max(nl, 1),
                    "cells": {"reference_equivalent": int(ctr["dp_cells"] // max(args.steps, 1)), "computed": int(ctr["dp_cells_computed"] // max(args.steps, 1)),
                              "gcups_on_computed_cells": fill_gcups_computed, "frac_on_computed_cells": fill_gcups_computed / alu_peak},
                    "how": "peak = 148 SMs x 1.965 GHz x %.2f ALU-pipe warp-inst/clk/SM (tools/pipe_bench.cu, profiles/r1_pipe_bench.txt) x 64 cells / %d ALU-pipe "
                           "instructions per cell pair; achieved = rdlen x rflen cells of the DP problems the reference's control flow runs / K_fill time"
                           % (ALU_WARP_INST_PER_CLK_SM, ALU_INST_PER_CELL_PAIR),
                    "hbm": hbm,
                    "k_bt": {"ms_per_step": phases["k_bt"][0] / args.steps, "bound": "issue under divergence / dependent 16-byte chunk loads (ncu: 16 of 32 lanes active, "
                             "issue 55 %): no bandwidth or ALU ceiling applies", "algorithmic_GBps": (phases["k_bt"][2] / 1e9) / (phases["k_bt"][0] * 1e-3) if phases["k_bt"][0] > 0 else None}}
    else:
        roofline = dict(hbm, bound="hbm", kernel=dom, traffic=traffic, launches=int(nl), ms_per_launch=ms / max(nl, 1))
    line = {"metric": METRIC, "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": workload_config(n), "aligned_fraction": aligned / float(n),
            "gcups": ctr["dp_cells"] * world / dev_s / 1e9 if dev_s > 0 else None,
            "gcups_fill_kernel": fill_gcups,
            "phase_ms_per_step": {k: phases[k][0] / args.steps for k in phases} | {"glue": ctr["ms_glue"] / args.steps, "wall": 1000.0 * wall / args.steps},
            "roofline": roofline,
            "int_alu": {"kernel": "k_fill", "achieved_gcups": fill_gcups, "peak_gcups": alu_peak, "frac": fill_gcups / alu_peak},
            "e2e": {"value": total_reads / e2e_s, "unit": "reads/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(ctr["launches"]), "clocks": clocks}
    if not args.no_file_to_file:
        try:
            if world == 1:
                line["file_to_file"] = file_to_file(pkg, job, ascii_reads, local)
            else:
                f2f = file_to_file_sharded(pkg, job, ascii_reads, local, rank, world, dist, torch)
                if rank == 0:
                    line["file_to_file"] = f2f
        except Exception as e:
            line["file_to_file"] = {"value": None, "error": str(e)[:200]}
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            try:   # bounded: a two-point -p sweep on 60 k reads, then one run on a 200 k-read prefix of the workload (~20 s of CPU work)
                cores = cpu_cores()
                threads, pm, sweep = tune_reference(base, codes, cores, quick=True)
                sample = int(os.environ.get("BT2_BENCH_CPU_READS", "200000"))
                fq = os.path.join(WORK, "ref_sample.fq")
                write_fastq(fq, ascii_reads[:sample])
                cb = summarize_reference([run_reference_binary(base, fq, sample, threads, pm)], sweep, cores)
                cb["value"] = cb["hot_path_reads_per_s"]
                cb["sample"] = ("first %d reads of the workload, bowtie2-align-l OMP_NUM_THREADS=%d -p %d -S /dev/null; value = reads covered by the reference's "
                                "phase timers / sum of its P1-P6 phase times (hot path only); wall_reads_per_s is the whole process" % (sample, threads, pm))
                line["cpu_baseline"] = cb
            except Exception as e:  # the baseline is reported, never required
                line["cpu_baseline"] = {"value": None, "unit": "reads/s", "cores": cpu_cores(), "kind": "reference", "sample": "failed: %s" % e}
        print(json.dumps(line), flush=True)
    gpu.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
