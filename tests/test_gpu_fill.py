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


def _pack(probs):
    n = len(probs)
    a = {"seq_off": np.zeros(n, np.uint32), "rf_off": np.zeros(n, np.uint32), "hef_off": np.zeros(n, np.uint64),
         "rows": np.array([p[3] for p in probs], np.uint16), "cols": np.array([p[4] for p in probs], np.uint16),
         "minsc": np.array([p[5] for p in probs], np.int32)}
    so = ro = ho = 0
    for i, p in enumerate(probs):
        a["seq_off"][i] = so; a["rf_off"][i] = ro; a["hef_off"][i] = ho
        so += p[3]; ro += p[4] + 1; ho += p[3] * p[4]
    a["seq"] = np.concatenate([p[0] for p in probs]); a["qual"] = np.concatenate([p[1] for p in probs]); a["rf"] = np.concatenate([p[2] for p in probs])
    a["ncells"] = ho
    return a


def run_test_fill(g, gpu, a, mode):
    """bt2gpu_test_fill: mode 0 = test instantiation over the whole rectangle (+ H/E/F), 1 = production instantiation over the band"""
    n, ho = len(a["rows"]), a["ncells"]
    out = {k: np.zeros(ho, np.uint8) for k in ("H", "E", "F", "FL")}
    out["cand"] = np.zeros(256 * n, np.uint16); out["ncand"] = np.zeros(n, np.uint32); out["lrmax"] = np.zeros(n, np.int32); out["band"] = np.zeros(2 * n, np.int32)
    ptr = lambda x: x.ctypes.data_as(C.c_void_p)
    g.bt2gpu_test_fill.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 16 + [C.c_int32, C.c_void_p]
    rc = g.bt2gpu_test_fill(gpu.ptr, n, ptr(a["seq_off"]), ptr(a["seq"]), ptr(a["qual"]), ptr(a["rows"]), ptr(a["cols"]), ptr(a["rf_off"]), ptr(a["rf"]),
                            ptr(a["minsc"]), ptr(a["hef_off"]), ptr(out["H"]), ptr(out["E"]), ptr(out["F"]), ptr(out["FL"]), ptr(out["cand"]),
                            ptr(out["ncand"]), ptr(out["lrmax"]), mode, ptr(out["band"]))
    assert rc == 0, g.bt2gpu_last_error(gpu.ptr)
    return out


def walked_cells(fl, cands, rows, cols, gapbar):
    """The cells SwAligner::nextAlignment's candidate loop + backtraceNucleotidesEnd2EndSseU8 stand on, from the per-cell comparison
    flags alone (the walk of csrc/gpu/dev_bt.cuh::hot: fixed priority diagonal > ref-gap open > ref-gap extend > read-gap open >
    read-gap extend, a cell already visited ends the attempt).  Every candidate is tried (a superset of what any policy tries)."""
    vis = set()
    for key in cands:
        row, col, ct = rows - 1, int(key) & 255, 2
        if (row, col) in vis:
            continue
        while True:
            if (row, col) in vis:
                break
            vis.add((row, col))
            f = ~int(fl[row, col]) & 0x7f
            left = col > 0
            eo = (((f >> 2) & 2) | ((f >> 4) & 1)) if left else 0
            fo = ((f >> 2) & 8) | ((f >> 4) & 4)
            ho = 16 if (left and (f & CF_D)) else 0
            if gapbar <= row < rows - gapbar:
                if f & CF_HF:
                    ho |= fo
                if f & CF_HE:
                    ho |= eo
            opts = eo if ct == 0 else (fo if ct == 1 else ho)
            if row == 0 or opts == 0:
                break
            sel = opts.bit_length() - 1
            ct = 1 if sel == 2 else (0 if sel == 0 else 2)
            if sel >= 2:
                row -= 1
            if sel not in (2, 3):
                col -= 1
    return vis


def test_fill_kernel_matches_oracle():
    """mode 0: every H/E/F byte, every (specified) flag bit, last-row maximum and candidates of the whole rectangle == the oracle's
    scalar recurrence.  mode 1 (what bt2gpu_align_batch runs): same candidates, and the same flags on every cell a backtrace can
    stand on; the cells it leaves out lie outside the band."""
    job = pkg.Job(common.index_base("lambda"), "")
    gpu = pkg.GpuAligner(job, 0)
    g = pkg.load_gpu()
    orc = C.CDLL(common.ROOT + "/oracle/liboracle.so")
    rng = np.random.default_rng(99)
    probs = _problems(rng, 300)
    a = _pack(probs)
    full = run_test_fill(g, gpu, a, 0)
    band = run_test_fill(g, gpu, a, 1)
    ptr = lambda x: x.ctypes.data_as(C.c_void_p)
    orc.oracle_fill.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6
    P = job.params.contents
    nwalked = ncomputed = ncells = 0
    for i, p in enumerate(probs):
        r, c = p[3], p[4]
        oH = np.zeros(r * c, np.uint8); oE = np.zeros(r * c, np.uint8); oF = np.zeros(r * c, np.uint8)
        oc = np.zeros(256, np.uint16); on = C.c_int(0); olr = C.c_int(0)
        orc.oracle_fill(C.cast(job.params, C.c_void_p), ptr(p[0]), ptr(p[1]), r, ptr(p[2]), c, p[5], ptr(oH), ptr(oE), ptr(oF), ptr(oc), C.byref(on), C.byref(olr))
        o = int(a["hef_off"][i])
        for nm, x, b in (("H", full["H"], oH), ("E", full["E"], oE), ("F", full["F"], oF)):
            got = x[o:o + r * c]
            if (got != b).any():
                k = int(np.nonzero(got != b)[0][0])
                raise AssertionError("problem %d (%dx%d) %s differs at row %d col %d: gpu %d oracle %d" % (i, r, c, nm, k // c, k % c, got[k], b[k]))
        ne, care = expected_flags(P, p[0], p[1], p[2], oH.reshape(r, c), oE.reshape(r, c), oF.reshape(r, c))
        got = full["FL"][o:o + r * c].reshape(r, c)
        bad = ((got ^ ne) & care) != 0
        if bad.any():
            rr, cc = [int(x[0]) for x in np.nonzero(bad)]
            raise AssertionError("problem %d (%dx%d) flags differ at row %d col %d: gpu %02x expected %02x (care %02x)" % (i, r, c, rr, cc, got[rr, cc], ne[rr, cc], care[rr, cc]))
        for res, what in ((full, "whole rectangle"), (band, "band")):
            if what == "band" and on.value == 0:
                assert res["ncand"][i] == 0 and res["lrmax"][i] <= olr.value, (i, what)     # no candidate: the band's maximum is a lower bound
                continue
            assert res["lrmax"][i] == olr.value, (i, what, res["lrmax"][i], olr.value)
            assert res["ncand"][i] == on.value, (i, what, res["ncand"][i], on.value)
            assert (res["cand"][256 * i:256 * i + on.value] == oc[:on.value]).all(), (i, what)
        # production band: whatever a backtrace can stand on is there and equal
        bfl = band["FL"][o:o + r * c].reshape(r, c)
        d0, nd = int(band["band"][2 * i]), int(band["band"][2 * i + 1])
        dd = np.arange(c)[None, :] - np.arange(r)[:, None] - d0
        inband = (dd >= 0) & (dd < nd)
        assert ((bfl == 0x80) == ~inband).all(), "problem %d: cells computed do not match the reported band" % i
        cells = walked_cells(got, oc[:on.value], r, c, P.gapbar)
        for (rr, cc) in cells:
            assert inband[rr, cc], "problem %d (%dx%d minsc %d): a backtrace reaches row %d col %d outside the band [%d, %d)" % (i, r, c, p[5], rr, cc, d0, d0 + nd)
            m = care[rr, cc] & 0x7f
            assert (bfl[rr, cc] & m) == (got[rr, cc] & m), "problem %d: band flags differ on a walked cell (row %d col %d): %02x vs %02x" % (i, rr, cc, bfl[rr, cc], got[rr, cc])
        nwalked += len(cells); ncomputed += int(inband.sum()); ncells += r * c
    print("fill: %d cells, %d in the bands computed, %d walked" % (ncells, ncomputed, nwalked))
    assert nwalked > 10000
    gpu.close()
