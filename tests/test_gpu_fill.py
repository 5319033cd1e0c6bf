"""GPU: the DP-fill kernel alone (bt2gpu_test_fill hook) against the oracle's scalar
recurrence: every H/E/F byte of the rectangle (debug instantiation), the byte of
backtrace-comparison flags the production kernel stores per cell (recomputed here from
the oracle's H/E/F with the integer comparisons of backtraceNucleotidesEnd2EndSseU8,
aligner_swsse_ee_u8.cpp:820-1100), the last-row maximum and the sorted candidate list,
for random reads/windows of ragged sizes (1..160 rows, 1..199 columns)."""
import ctypes as C
import numpy as np
import pytest
import common
from common import pkg

pytestmark = pytest.mark.gpu

CF_D, CF_HE, CF_HF, CF_EO, CF_EX, CF_FO, CF_FX, CF_VIS = 1, 2, 4, 8, 16, 32, 64, 128   # CF_VIS: written 0 by the fill (the backtrace's visited bit)


def expected_flags(P, rd, qual, masks, oH, oE, oF):
    """"not equal" flag byte per cell + a byte mask of the bits that are specified
    (DESIGN.md "cell flags": column 0 has no left neighbour, row 0 is never examined, E/F
    comparisons of rows inside the gap barrier are unreachable)."""
    r, c = oH.shape
    H = oH.astype(np.int32); E = oE.astype(np.int32); F = oF.astype(np.int32)
    rdgo = P.rdGapConst + P.rdGapLinear; rdge = P.rdGapLinear; rfgo = P.rfGapConst + P.rfGapLinear; rfge = P.rfGapLinear
    mmp = np.array(P.mmpens[:], np.int32); npn = np.array(P.npens[:], np.int32)
    q = np.clip(qual.astype(np.int32) - 33, 0, 255)
    refc = np.where(masks[:c] > 15, 4, np.log2(np.maximum(masks[:c], 1)).astype(np.int32))
    isn = (rd[:, None] > 3) | (refc[None, :] == 4)
    pen = np.where(isn, npn[q][:, None], np.where(rd[:, None] == refc[None, :], 0, mmp[q][:, None]))
    Hd = np.zeros((r, c), np.int32); Hd[0, :] = 255; Hd[1:, 1:] = H[:-1, :-1]          # H(r-1,c-1); H(-1,*) = 0xff, H(*,-1) = 0
    Hl = np.zeros((r, c), np.int32); Hl[:, 1:] = H[:, :-1]
    El = np.zeros((r, c), np.int32); El[:, 1:] = E[:, :-1]
    Hu = np.full((r, c), 255, np.int32); Hu[1:, :] = H[:-1, :]
    Fu = np.zeros((r, c), np.int32); Fu[1:, :] = F[:-1, :]
    ne = np.zeros((r, c), np.uint8)
    ne |= np.where(H != Hd - pen, CF_D, 0).astype(np.uint8)
    ne |= np.where(H != E, CF_HE, 0).astype(np.uint8)
    ne |= np.where(H != F, CF_HF, 0).astype(np.uint8)
    ne |= np.where(E != Hl - rdgo, CF_EO, 0).astype(np.uint8)
    ne |= np.where(E != El - rdge, CF_EX, 0).astype(np.uint8)
    ne |= np.where(F != Hu - rfgo, CF_FO, 0).astype(np.uint8)
    ne |= np.where(F != Fu - rfge, CF_FX, 0).astype(np.uint8)
    care = np.full((r, c), 255, np.uint8)
    care[:, 0] &= ~np.uint8(CF_EO | CF_EX)
    rows_i = np.arange(r)
    barrier = (rows_i < P.gapbar) | (r - 1 - rows_i < P.gapbar)
    care[barrier, :] &= ~np.uint8(CF_EO | CF_EX | CF_FO | CF_FX | CF_HE | CF_HF)
    care[0, :] = CF_VIS | CF_D
    return ne, care


def _problems(rng, n):
    probs = []
    for i in range(n):
        rows = int(rng.integers(1, 161)) if i > 8 else [1, 2, 63, 64, 65, 128, 129, 150, 160][i]
        cols = int(rng.integers(1, 200)) if i > 8 else [1, 5, 199, 100, 70, 199, 150, 199, 199][i]
        ref = rng.integers(0, 4, size=cols + 1).astype(np.uint8)
        # the read is a mutated copy of a window so that real alignments exist
        st = int(rng.integers(0, max(1, cols - rows + 1)))
        rd = np.resize(ref[st:], rows).copy()
        mut = rng.random(rows) < 0.05
        rd[mut] = rng.integers(0, 4, size=int(mut.sum()))
        if rows > 30 and rng.random() < 0.5:
            p = int(rng.integers(5, rows - 5)); rd = np.concatenate([rd[:p], rd[p + 1:], rd[:1]])
        if rng.random() < 0.3:
            rd[rng.integers(0, rows)] = 4
        masks = (1 << ref).astype(np.uint8)
        if rng.random() < 0.3:
            masks[rng.integers(0, cols)] = 16
        qual = rng.integers(35, 74, size=rows).astype(np.uint8)
        minsc = -int(rng.integers(0, 120))
        probs.append((rd.astype(np.uint8), qual, masks, rows, cols, minsc))
    return probs


def test_fill_kernel_matches_oracle():
    job = pkg.Job(common.index_base("lambda"), "")
    gpu = pkg.GpuAligner(job, 0)
    g = pkg.load_gpu()
    orc = C.CDLL(common.ROOT + "/oracle/liboracle.so")
    rng = np.random.default_rng(99)
    probs = _problems(rng, 300)
    n = len(probs)
    seq_off = np.zeros(n, np.uint32); rf_off = np.zeros(n, np.uint32); hef_off = np.zeros(n, np.uint64)
    rows = np.array([p[3] for p in probs], np.uint16); cols = np.array([p[4] for p in probs], np.uint16)
    minsc = np.array([p[5] for p in probs], np.int32)
    so = ro = ho = 0
    for i, p in enumerate(probs):
        seq_off[i] = so; rf_off[i] = ro; hef_off[i] = ho
        so += p[3]; ro += p[4] + 1; ho += p[3] * p[4]
    seq = np.concatenate([p[0] for p in probs]); qual = np.concatenate([p[1] for p in probs]); rf = np.concatenate([p[2] for p in probs])
    H = np.zeros(ho, np.uint8); E = np.zeros(ho, np.uint8); F = np.zeros(ho, np.uint8); FL = np.zeros(ho, np.uint8)
    cand = np.zeros(256 * n, np.uint16); ncand = np.zeros(n, np.uint32); lrmax = np.zeros(n, np.int32)
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    g.bt2gpu_test_fill.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 16
    rc = g.bt2gpu_test_fill(gpu.ptr, n, ptr(seq_off), ptr(seq), ptr(qual), ptr(rows), ptr(cols), ptr(rf_off), ptr(rf), ptr(minsc),
                            ptr(hef_off), ptr(H), ptr(E), ptr(F), ptr(FL), ptr(cand), ptr(ncand), ptr(lrmax))
    assert rc == 0, g.bt2gpu_last_error(gpu.ptr)
    orc.oracle_fill.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6
    for i, p in enumerate(probs):
        r, c = p[3], p[4]
        oH = np.zeros(r * c, np.uint8); oE = np.zeros(r * c, np.uint8); oF = np.zeros(r * c, np.uint8)
        oc = np.zeros(256, np.uint16); on = C.c_int(0); olr = C.c_int(0)
        orc.oracle_fill(C.cast(job.params, C.c_void_p), ptr(p[0]), ptr(p[1]), r, ptr(p[2]), c, p[5], ptr(oH), ptr(oE), ptr(oF), ptr(oc), C.byref(on), C.byref(olr))
        o = int(hef_off[i])
        for nm, a, b in (("H", H, oH), ("E", E, oE), ("F", F, oF)):
            got = a[o:o + r * c]
            if (got != b).any():
                k = int(np.nonzero(got != b)[0][0])
                raise AssertionError("problem %d (%dx%d) %s differs at row %d col %d: gpu %d oracle %d" % (i, r, c, nm, k // c, k % c, got[k], b[k]))
        ne, care = expected_flags(job.params.contents, p[0], p[1], p[2], oH.reshape(r, c), oE.reshape(r, c), oF.reshape(r, c))
        got = FL[o:o + r * c].reshape(r, c)
        bad = ((got ^ ne) & care) != 0
        if bad.any():
            rr, cc = [int(x[0]) for x in np.nonzero(bad)]
            raise AssertionError("problem %d (%dx%d) flags differ at row %d col %d: gpu %02x expected %02x (care %02x)" % (i, r, c, rr, cc, got[rr, cc], ne[rr, cc], care[rr, cc]))
        assert lrmax[i] == olr.value, (i, lrmax[i], olr.value)
        assert ncand[i] == on.value, (i, ncand[i], on.value)
        assert (cand[256 * i:256 * i + on.value] == oc[:on.value]).all(), i
    gpu.close()
