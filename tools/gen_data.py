#!/usr/bin/env python
"""Deterministic synthetic reference + simulated reads (SURVEY 8(d) C2-C4 restated; the
reference's own scripts/sim needs Math::Random, absent here).
  gen_data.py ref  OUT.fa  LEN  [--seed S] [--repeats FRAC] [--nfam F] [--nseq K] [--ngaps G]
  gen_data.py reads REF.fa OUT.fq N LEN [--seed S] [--subst 0.01] [--indel 0.10] [--randqual] [--nfrac 0.0]
"""
import sys, argparse
import numpy as np

BASES = np.frombuffer(b'ACGT', dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8); COMP[list(b'ACGTN')] = list(b'TGCAN')

def gen_ref(a):
    rng = np.random.default_rng(a.seed)
    codes = rng.integers(0, 4, size=a.len, dtype=np.uint8)
    if a.repeats > 0:
        # repeat families: 50 families x 300 bp, copies with 2% divergence until FRAC of the text is covered
        nfam, flen = a.nfam, 300
        fams = rng.integers(0, 4, size=(nfam, flen), dtype=np.uint8)
        ncopies = int(a.repeats * a.len / flen)
        pos = rng.integers(0, a.len - flen, size=ncopies)
        for p in pos:
            f = fams[rng.integers(0, nfam)].copy()
            m = rng.random(flen) < 0.02
            f[m] = (f[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) % 4
            codes[p:p + flen] = f
    seq = BASES[codes]
    if a.ngaps > 0:
        # runs of N (1..150 bp) at random places, one of them at the very start and one at the very end of the text
        seq = seq.copy()
        starts = rng.integers(0, a.len - 200, size=a.ngaps); lens = rng.integers(1, 151, size=a.ngaps)
        starts[0] = 0; starts[-1] = a.len - int(lens[-1])
        for st, ln in zip(starts, lens): seq[st:st + ln] = ord('N')
    with open(a.out, 'wb') as f:
        per = a.len // a.nseq
        for k in range(a.nseq):
            s = seq[k * per:(k + 1) * per if k < a.nseq - 1 else a.len]
            f.write(b'>seq%d\n' % k)
            for i in range(0, len(s), 60): f.write(s[i:i + 60].tobytes() + b'\n')

def load_fa(path):
    seqs = []; cur = []
    for ln in open(path, 'rb'):
        if ln.startswith(b'>'):
            if cur: seqs.append(b''.join(cur)); cur = []
        else: cur.append(ln.strip())
    if cur: seqs.append(b''.join(cur))
    return [np.frombuffer(s, dtype=np.uint8) for s in seqs]

def gen_reads(a):
    rng = np.random.default_rng(a.seed)
    seqs = load_fa(a.ref)
    out = open(a.out, 'wb')
    for i in range(a.n):
        s = seqs[rng.integers(0, len(seqs))]
        L = a.len if a.maxlen <= a.len else int(rng.integers(a.len, a.maxlen + 1))
        st = int(rng.integers(0, len(s) - L - 2))
        r = s[st:st + L + 2].copy()
        m = rng.random(L + 2) < a.subst
        if m.any():
            idx = np.searchsorted(BASES, r[m])
            r[m] = BASES[(idx + rng.integers(1, 4, size=int(m.sum()))) % 4]
        if rng.random() < a.indel and L > 24:
            p = int(rng.integers(10, L - 10))
            if rng.random() < 0.5: r = np.concatenate([r[:p], BASES[rng.integers(0, 4, size=1)], r[p:]])  # insertion in read
            else: r = np.concatenate([r[:p], r[p + 1:]])                                                 # deletion
        r = r[:L]
        if a.nfrac > 0:
            nm = rng.random(L) < a.nfrac
            r[nm] = ord('N')
        if rng.random() < 0.5: r = COMP[r[::-1]]
        if a.randqual: q = rng.integers(35, 74, size=L).astype(np.uint8)   # Q2..Q40
        else: q = np.full(L, ord('I'), dtype=np.uint8)
        out.write(b'@r%d\n' % i + r.tobytes() + b'\n+\n' + q.tobytes() + b'\n')
    out.close()

def main():
    ap = argparse.ArgumentParser(); sub = ap.add_subparsers(dest='cmd')
    p = sub.add_parser('ref'); p.add_argument('out'); p.add_argument('len', type=int); p.add_argument('--seed', type=int, default=12345)
    p.add_argument('--repeats', type=float, default=0.0); p.add_argument('--nfam', type=int, default=50); p.add_argument('--nseq', type=int, default=1); p.add_argument('--ngaps', type=int, default=0)
    p = sub.add_parser('reads'); p.add_argument('ref'); p.add_argument('out'); p.add_argument('n', type=int); p.add_argument('len', type=int)
    p.add_argument('--seed', type=int, default=2); p.add_argument('--subst', type=float, default=0.01); p.add_argument('--indel', type=float, default=0.10)
    p.add_argument('--randqual', action='store_true'); p.add_argument('--nfrac', type=float, default=0.0); p.add_argument('--maxlen', type=int, default=0)
    a = ap.parse_args()
    gen_ref(a) if a.cmd == 'ref' else gen_reads(a)

if __name__ == '__main__':
    main()
