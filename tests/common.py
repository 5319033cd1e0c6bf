"""Shared helpers of the test-suite: case table, index cache, oracle binding.
The oracle (oracle/liboracle.so) is TEST INFRASTRUCTURE: it is only ever loaded from here,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
WORK = os.path.join(ROOT, "tests", "_work")
REFBIN = os.path.join(ROOT, "oracle", "_ref")
sys.path.insert(0, ROOT)
pkg = importlib.import_module("omp-bowtie2-prime_b200")

# name -> (reference fasta maker, reads maker, options, golden file)
CASES = {
    "lambda_default": ("lambda", "lambda_reads_1k", "", "lambda_default.sam"),
    "lambda_k3_vs": ("lambda", "lambda_reads_1k", "-k 3 --very-sensitive", "lambda_k3_vs.sam"),
    "syn150_sensitive_dpad12": ("syn", "syn150", "--sensitive --dpad 12", "syn150_sensitive_dpad12.sam"),
    "syn100_default": ("syn", "syn100", "", "syn100_default.sam"),
    "syn100_k5": ("syn", "syn100", "-k 5", "syn100_k5.sam"),
    "synN120_default": ("synN", "synN120", "", "synN120_default.sam"),
    # high-copy repeats: seed ranges of 128+ rows (row sampling in "seen list" mode and its conversion), -D streaks, several DPs per read
    "synR100_default": ("synR", "synR100", "", "synR100_default.sam"),
    "synR100_k3_vs": ("synR", "synR100", "-k 3 --very-sensitive", "synR100_k3_vs.sam"),
    # the reference built with -DENABLE_I16 and run with --no-sse8: its 16-bit DP path (other RNG protocol)
    "syn100_k5_i16": ("syn", "syn100", "-k 5 --i16 --no-sse8", "syn100_k5_i16.sam"),
    "syn150_sens_i16": ("syn", "syn150", "--i16 --no-sse8 --sensitive --dpad 12", "syn150_sens_i16.sam"),
    # minimum score -300 < -254: that build takes the 16-bit path with 8-bit SSE still enabled; the score range a8-bit fill does not have
    "syn100_min300_i16": ("syn", "syn100", "--i16 --score-min L,-300,0 -k 3", "syn100_min300_i16.sam"),
}


# Reads longer than the reference's 160-row limit: the golden SAM comes from oracle/_ref/bowtie2-align-l-long, the
# -DENABLE_I16 reference with ALN_MAX_ROWS / ALN_MAX_COLS raised in a scratch copy (oracle/build_ref.sh) -- not the reference as shipped.
# 161..255 bp is where that binary is a coherent oracle: its seed layer keeps the read length and the seed offsets in 8-bit fields
# (aligner_seed.h:679-683, 1226-1230), so beyond 255 bp it cuts seeds from (offset mod 256) and frames them on the wrong diagonal.  The
# restatement and the CUDA path use the offsets as they are; on the 250/600/1000 bp set below the SAM still comes out the same (the
# wrapped seeds only add failed extensions), at scale ~1 % of such reads differ.
CASES_LONG = {
    "synL255_i16": ("syn", "synL255", "--i16", "synL255_i16.sam"),
    "synL255_k3_sens": ("syn", "synL255", "--i16 --sensitive -k 3", "synL255_k3_sens.sam"),
    "synL_mix_i16": ("syn", "synLmix", "--i16", "synL_mix_i16.sam"),
    "synL_mix_k3_sens": ("syn", "synLmix", "--i16 --sensitive -k 3", "synL_mix_k3_sens.sam"),
}
LONG_SETS = {"synLmix": ((250, 150, 21, 0.01), (600, 120, 22, 0.02), (1000, 120, 23, 0.02)),
             "synL255": ((170, 400, 31, 0.02), (200, 400, 32, 0.02), (230, 400, 33, 0.03), (255, 400, 34, 0.02))}


def _run(cmd):
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def have_builder():
    return os.access(os.path.join(REFBIN, "bowtie2-build-l"), os.X_OK)


def gen(args):
    _run([sys.executable, os.path.join(ROOT, "tools", "gen_data.py")] + [str(a) for a in args])


def index_base(name):
    """Build (once) the .bt2l index of a named reference with the reference's own builder."""
    os.makedirs(WORK, exist_ok=True)
    base = os.path.join(WORK, name)
    if os.path.exists(base + ".4.bt2l") and os.path.exists(base + ".1.bt2l"):
        return base
    if name == "lambda":
        fa = os.path.join(GOLD, "lambda_virus.fa")
    elif name == "syn":
        fa = base + ".fa"
        gen(["ref", fa, 300000, "--seed", 4242, "--repeats", 0.2, "--nseq", 3])
    elif name == "synN":
        fa = base + ".fa"
        gen(["ref", fa, 200000, "--seed", 99, "--repeats", 0.1, "--nseq", 2, "--ngaps", 40])
    elif name == "synR":
        fa = base + ".fa"
        gen(["ref", fa, 200000, "--seed", 77, "--repeats", 0.45, "--nfam", 2])
    elif name.startswith("iid"):  # iid<len>
        fa = base + ".fa"
        gen(["ref", fa, int(name[3:]), "--seed", 12345])
    else:
        raise KeyError(name)
    _run([os.path.join(REFBIN, "bowtie2-build-l"), "--threads", "4", fa, base])
    return base


def reads_path(name):
    os.makedirs(WORK, exist_ok=True)
    if name == "lambda_reads_1k":
        return os.path.join(GOLD, "lambda_reads_1k.fq")
    p = os.path.join(WORK, name + ".fq")
    if os.path.exists(p):
        return p
    if name in LONG_SETS:                          # reads of several lengths, names made unique per length
        index_base("syn")
        with open(p + ".tmp", "w") as out:
            for L, nr, seed, subst in LONG_SETS[name]:
                part = os.path.join(WORK, "synL%d.fq" % L)
                gen(["reads", os.path.join(WORK, "syn.fa"), part, nr, L, "--seed", seed, "--subst", subst])
                for i, line in enumerate(open(part)):
                    out.write("@L%d_%s" % (L, line[1:]) if i % 4 == 0 else line)
        os.replace(p + ".tmp", p)
        return p
    if name == "synN120":
        index_base("synN")
        gen(["reads", os.path.join(WORK, "synN.fa"), p, 1500, 120, "--seed", 9, "--randqual"])
        return p
    if name == "synR100":
        index_base("synR")
        gen(["reads", os.path.join(WORK, "synR.fa"), p, 1500, 100, "--seed", 11, "--subst", 0.02])
        return p
    index_base("syn")
    fa = os.path.join(WORK, "syn.fa")
    if name == "syn150":
        gen(["reads", fa, p, 1500, 150, "--seed", 7, "--randqual", "--nfrac", 0.002])
    elif name == "syn100":
        gen(["reads", fa, p, 1500, 100, "--seed", 8, "--subst", 0.03])
    else:
        raise KeyError(name)
    return p


def sorted_records(sam_text):
    recs = [l for l in sam_text.split("\n") if l and not l.startswith("@")]
    return sorted(recs, key=lambda s: s.encode("latin-1"))


def golden_records(fname):
    with open(os.path.join(GOLD, fname), encoding="latin-1") as f:
        return [l.rstrip("\n") for l in f if l.strip()]


class Oracle:
    """ctypes view of oracle/liboracle.so (oracle_capi.cpp); long_reads: liboracle_long.so, the restatement built with the reference's
    ALN_MAX_ROWS / ALN_MAX_COLS raised to 1000 / 1072 (oracle D: beyond the reference's limits)."""

    def __init__(self, job, long_reads=False):
        p = os.path.join(ROOT, "oracle", "liboracle_long.so" if long_reads else "liboracle.so")
        self.lib = C.CDLL(p)
        self.lib.oracle_create.restype = C.c_void_p
        self.lib.oracle_create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.oracle_align_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        self.lib.oracle_destroy.argtypes = [C.c_void_p]
        self.lib.oracle_counters.argtypes = [C.c_void_p, C.POINTER(C.c_uint64 * 10)]
        self.ptr = self.lib.oracle_create(C.cast(job.index, C.c_void_p), C.cast(job.ref, C.c_void_p), C.cast(job.params, C.c_void_p))
        self.job = job

    def align(self, batch, nthreads=0):
        r = pkg.bt2gpu_results()
        self.lib.oracle_align_batch(self.ptr, C.cast(batch.reads, C.c_void_p), C.cast(C.byref(r), C.c_void_p), nthreads)
        return pkg.Results(r)

    def counters(self):
        a = (C.c_uint64 * 10)()
        self.lib.oracle_counters(self.ptr, C.byref(a))
        names = ["seeds", "lfSearch", "lfExtend", "saRows", "lfResolve", "dps", "cells", "btAttempts", "btSteps", "alns"]
        return dict(zip(names, list(a)))

    def close(self):
        if self.ptr:
            self.lib.oracle_destroy(self.ptr)
            self.ptr = None


def compare_results(a, b, check_counters=True, i16path=False, exact_xs=False):
    """Field-for-field comparison of two Results (oracle vs CUDA). Returns a list of differences.
    prm.bestLtMinscMate1 (the best score of a read's windows that had NO valid alignment; reaches SAM only as Xs:i / ZT:Z under
    --sam-opt-config): K_fill computes the band of diagonals valid alignments can lie in, so b's value is a lower bound of a's
    unless the job runs with --exact-xs (bt2gpu_params.exactBestLtMinsc), where it is compared exactly.
    i16path: prm.bestLtMinscMate1 is left out -- under the 16-bit path the reference stores `raw 16-bit word - 0xff` there
    (aligner_swsse_ee_i16.cpp:640), which needs the unsaturated score of windows that fail; the CUDA fill keeps 8-bit saturation."""
    import numpy as np
    diffs = []
    if len(a.reads) != len(b.reads):
        return ["n reads %d vs %d" % (len(a.reads), len(b.reads))]
    fields = ["n_alns", "rng_last", "rng_lastOff", "minsc", "status", "flags", "rounds", "done_policy"]
    if check_counters:
        fields += ["nExIters", "nExDps", "nExDpSuccs", "nExDpFails", "nDpFail", "nDpFailStreak", "nDpLastSucc", "nRedundants",
                   "bestLtMinsc", "nonzTot", "numElts", "numRanges", "nSeedsTried", "n_scratch"]
    if "bestLtMinsc" in fields and (i16path or not exact_xs):
        fields.remove("bestLtMinsc")
        if not i16path:
            lo = np.iinfo(np.int32).min
            bad = np.nonzero((b.reads["bestLtMinsc"] > a.reads["bestLtMinsc"]) | ((a.reads["bestLtMinsc"] == lo) != (b.reads["bestLtMinsc"] == lo)))[0]
            if len(bad):
                diffs.append("reads.bestLtMinsc is not a lower bound for %d reads, first %d: %s vs %s" % (len(bad), bad[0], a.reads["bestLtMinsc"][bad[0]], b.reads["bestLtMinsc"][bad[0]]))
    for f in fields:
        bad = np.nonzero(a.reads[f] != b.reads[f])[0]
        if len(bad):
            diffs.append("reads.%s differs for %d reads, first %d: %s vs %s" % (f, len(bad), bad[0], a.reads[f][bad[0]], b.reads[f][bad[0]]))
    if diffs:
        return diffs
    for i in range(len(a.reads)):
        ra, rb = a.reads[i], b.reads[i]
        for k in range(int(ra["n_alns"])):
            x, y = a.alns[ra["aln_off"] + k], b.alns[rb["aln_off"] + k]
            for f in ("refoff", "tidx", "score", "minsc", "ns", "gaps", "edits", "basesAligned", "refns", "n_edits", "fw"):
                if x[f] != y[f]:
                    diffs.append("read %d aln %d %s: %s vs %s" % (i, k, f, x[f], y[f]))
            ea = a.edits[x["edit_off"]:x["edit_off"] + x["n_edits"]]
            eb = b.edits[y["edit_off"]:y["edit_off"] + y["n_edits"]]
            if len(ea) != len(eb) or (ea != eb).any():
                diffs.append("read %d aln %d edits differ" % (i, k))
        sa = a.edits[ra["scratch_off"]:ra["scratch_off"] + ra["n_scratch"]]
        sb = b.edits[rb["scratch_off"]:rb["scratch_off"] + rb["n_scratch"]]
        if len(sa) != len(sb) or (sa != sb).any():
            diffs.append("read %d scratch differs" % i)
        if len(diffs) > 20:
            break
    return diffs
