#!/usr/bin/env python3
"""Second half of import_kat.sh: from the JSON dump of scripts/test/simple_tests.pl's case table keep what this fork can do at
all (unpaired reads given inline, end-to-end), translate the long-form options the table uses (--policy "A=..;B=..", --multiseed=...)
into the equivalent plain options -- checked here by running the UNMODIFIED reference binary with both spellings -- run that binary
on every case and store, per case: the reference/reads/options, the authors' expectation (hits, cigar, SAM optional flags), whether
the fork still meets it (it ignores -a, has exact seeds only, prints CIGAR/MD through an aliased edit buffer: SURVEY F3, F5, F14),
and the fork's own SAM records (the parity target).  usage: filter_kat.py all.json cases.json"""
import json
import os
import re
import shlex
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
REFBIN = os.path.join(ROOT, "oracle", "_ref")
SKIP_KEYS = ("mate1s", "mate2s", "pairhits", "paired", "tabbed", "fastq", "fasta", "qseq", "raw", "cline_reads", "cont_fasta_reads",
             "fastq1", "fasta1", "raw1", "qseq1", "cline_reads1", "should_abort", "idx_map", "lines", "names", "check_random")
BAD_OPTS = ("--local", "--trim-to", "-I ", "-X ", "--ff", "--fr", "--rf", "--no-mixed", "--no-discordant", "--un", "--al", "-f ", "-s ", "-u ",
            "--preserve-tags", "--xeq", "--sam-append-comment", "--sam-omit", "--align-paired-reads", "-F ", "--ma", "-3 ", "-5 ", "--n-ceil")


def translate(args):
    """long-form options -> plain ones; None when something has no plain spelling"""
    out = []
    toks = shlex.split(args.replace("\\;", ";"))
    i = 0
    while i < len(toks):
        t = toks[i]
        if "=" in t and t.startswith("--") and not t.startswith("--policy"):
            k, v = t.split("=", 1); toks[i:i + 1] = [k, v]; continue
        if t == "--policy":
            for kv in toks[i + 1].split(";"):
                if not kv:
                    continue
                k, v = kv.split("=", 1)
                if k == "MMP" and v.startswith("C"):
                    out += ["--mp", "%s,%s" % (v[1:], v[1:])]
                elif k == "MMP" and v.startswith("Q"):
                    p = v.split(","); out += ["--mp", ",".join(p[1:])] if len(p) > 1 else []
                elif k == "NP" and v.startswith("C"):
                    out += ["--np", v[1:]]
                elif k == "RDG":
                    out += ["--rdg", v]
                elif k == "RFG":
                    out += ["--rfg", v]
                elif k == "MIN":
                    out += ["--score-min", v]
                elif k == "SEED":
                    p = v.split(","); out += ["-N", p[0]]
                    if len(p) > 1:
                        out += ["-L", p[1]]
                elif k == "SEEDLEN":
                    out += ["-L", v]
                elif k == "IVAL":
                    out += ["-i", v]
                elif k == "DPS":
                    out += ["-D", v]
                elif k == "ROUNDS":
                    out += ["-R", v]
                elif k == "MA" and v == "0":
                    pass
                else:
                    return None
            i += 2; continue
        if t == "--multiseed":
            p = toks[i + 1].split(","); out += ["-N", p[0]]
            if len(p) > 1:
                out += ["-L", p[1]]
            if len(p) > 2:
                out += ["-i", ",".join(p[2:])]
            i += 2; continue
        if t in ("-a", "--quiet", "--end-to-end"):
            i += 1; continue
        out.append(t); i += 1
    return out


def run_ref(base, fq, opts):
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([os.path.join(REFBIN, "bowtie2-align-l"), "-x", base, "-U", fq, "-p", "4", "--no-hd"] + opts, env=env, capture_output=True)
    if r.returncode != 0:
        return None
    return sorted(l for l in r.stdout.decode("latin-1").split("\n") if l and not l.startswith("pad\t"))


def main():
    cases = json.load(open(sys.argv[1]))
    keep, stats = [], {"total": len(cases), "not_unpaired_inline": 0, "option_outside_the_fork": 0, "untranslatable": 0, "reference_rejects": 0, "spellings_differ": 0}
    tmp = tempfile.mkdtemp()
    for c in cases:
        if "reads" not in c or any(k in c for k in SKIP_KEYS):
            stats["not_unpaired_inline"] += 1; continue
        args = (c.get("args") or "") + " " + (c["report"] if c.get("report") is not None else "-a")
        if any(b in args + " " for b in BAD_OPTS):
            stats["option_outside_the_fork"] += 1; continue
        plain = translate(args)
        if plain is None:
            stats["untranslatable"] += 1; continue
        if c.get("norc"):
            plain.append("--norc")
        if c.get("nofw"):
            plain.append("--nofw")
        fa = os.path.join(tmp, "r.fa"); fq = os.path.join(tmp, "r.fq"); base = os.path.join(tmp, "idx")
        with open(fa, "w") as f:
            for i, s in enumerate(c["ref"]):
                f.write(">%d\n%s\n" % (i, s))
        quals = c.get("quals") or ["I" * len(r) for r in c["reads"]]
        with open(fq, "w") as f:
            for i, (r, q) in enumerate(zip(c["reads"], quals)):
                f.write("@%d\n%s\n+\n%s\n" % (i, r, q))
            if len(c["reads"]) % 16 == 1:
                # the fork loses the last read of a file whose read count is 1 modulo 16 (measured: 1 read -> 0 records, 17 -> 16,
                # 33 -> 32; its per-slot read-ahead buffers, pat.h:1283-1440): a padding read keeps the case's reads out of that corner
                f.write("@pad\n%s\n+\n%s\n" % (c["reads"][-1], quals[-1]))
        for e in (".1.bt2l", ".2.bt2l", ".3.bt2l", ".4.bt2l", ".rev.1.bt2l", ".rev.2.bt2l"):
            if os.path.exists(base + e):
                os.remove(base + e)
        b = subprocess.run([os.path.join(REFBIN, "bowtie2-build-l"), fa, base], capture_output=True)
        if b.returncode != 0:
            stats["reference_rejects"] += 1; continue
        orig = [t for t in shlex.split(args.replace("\\;", ";")) if t != "--quiet"]
        if c.get("norc"):
            orig.append("--norc")
        if c.get("nofw"):
            orig.append("--nofw")
        sam_orig, sam = run_ref(base, fq, orig), run_ref(base, fq, plain)
        if sam is None or sam_orig is None:
            stats["reference_rejects"] += 1; continue
        if sam != sam_orig:
            stats["spellings_differ"] += 1; continue
        if len(sam) != len(c["reads"]):
            stats["reference_rejects"] += 1; continue
        # does the fork still meet the authors' expectation on positions?
        meets = True
        for rec in sam:
            f = rec.split("\t"); i = int(f[0])
            if i >= len(c.get("hits", [])):
                continue
            exp = c["hits"][i]
            got = "*" if int(f[1]) & 4 else str(int(f[3]) - 1)
            if got not in exp:
                meets = False
        keep.append({"name": c.get("name", "case %d" % len(keep)), "ref": c["ref"], "reads": c["reads"], "quals": quals, "args": " ".join(plain), "args_as_written": args.strip(),
                     "expect": {k: c[k] for k in ("hits", "hits_are_superset", "cigar", "samoptflags") if k in c},
                     "fork_meets_expected_positions": meets, "reference_sam": sam})
    json.dump({"source": "scripts/test/simple_tests.pl:66-4424 of sfiligoi/omp-bowtie2-prime @474d07aa", "selection": stats, "cases": keep},
              open(sys.argv[2], "w"), indent=1, sort_keys=True)
    print(stats, "kept", len(keep), "of which the fork meets the expected positions:", sum(k["fork_meets_expected_positions"] for k in keep))


if __name__ == "__main__":
    main()
