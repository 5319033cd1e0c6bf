#!/usr/bin/env python
"""Compare two SAM files record-for-record after sorting (output order depends on the
thread schedule).  Reports per-field mismatch counts.
usage: samdiff.py A.sam B.sam [--ignore YF,MD,CIGAR]"""
import sys, collections

def load(path):
    recs = collections.defaultdict(list)
    for ln in open(path):
        if ln.startswith('@'): continue
        f = ln.rstrip('\n').split('\t')
        recs[f[0]].append(f)
    for k in recs: recs[k].sort()
    return recs

def fields(f):
    d = {'FLAG': f[1], 'RNAME': f[2], 'POS': f[3], 'MAPQ': f[4], 'CIGAR': f[5], 'SEQ': f[9], 'QUAL': f[10]}
    for t in f[11:]:
        d[t[:2]] = t
    return d

def main():
    a, b = load(sys.argv[1]), load(sys.argv[2])
    ign = set()
    if '--ignore' in sys.argv: ign = set(sys.argv[sys.argv.index('--ignore') + 1].split(','))
    bad = collections.Counter(); nrec = 0; shown = 0
    for k in sorted(set(a) | set(b)):
        ra, rb = a.get(k, []), b.get(k, [])
        if len(ra) != len(rb):
            bad['NRECORDS'] += 1
            if shown < 10: print('NREC', k, len(ra), len(rb)); shown += 1
            continue
        for x, y in zip(ra, rb):
            nrec += 1
            dx, dy = fields(x), fields(y)
            for key in sorted(set(dx) | set(dy)):
                if key in ign: continue
                if dx.get(key) != dy.get(key):
                    bad[key] += 1
                    if shown < 15 and key not in ('SEQ', 'QUAL'):
                        print('DIFF', k, key, dx.get(key), '|', dy.get(key)); shown += 1
    print('records', nrec, 'mismatches', dict(bad))
    return 1 if bad else 0

if __name__ == '__main__':
    sys.exit(main())
