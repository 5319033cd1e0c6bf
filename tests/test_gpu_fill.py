"""GPU: the DP-fill kernel alone (bt2gpu_test_fill hook) against the oracle's scalar
recurrence: every H/E/F byte of the rectangle, the last-row maximum and the sorted
candidate list, for random reads/windows of ragged sizes (1..160 rows, 1..199 columns)."""
import ctypes as C
import numpy as np
import pytest
import common
from common import pkg

pytestmark = pytest.mark.gpu


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
    H = np.zeros(ho, np.uint8); E = np.zeros(ho, np.uint8); F = np.zeros(ho, np.uint8)
    cand = np.zeros(256 * n, np.uint16); ncand = np.zeros(n, np.uint32); lrmax = np.zeros(n, np.int32)
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    g.bt2gpu_test_fill.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 15
    rc = g.bt2gpu_test_fill(gpu.ptr, n, ptr(seq_off), ptr(seq), ptr(qual), ptr(rows), ptr(cols), ptr(rf_off), ptr(rf), ptr(minsc),
                            ptr(hef_off), ptr(H), ptr(E), ptr(F), ptr(cand), ptr(ncand), ptr(lrmax))
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
        assert lrmax[i] == olr.value, (i, lrmax[i], olr.value)
        assert ncand[i] == on.value, (i, ncand[i], on.value)
        assert (cand[256 * i:256 * i + on.value] == oc[:on.value]).all(), i
    gpu.close()
