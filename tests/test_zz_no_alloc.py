"""GPU: bt2gpu_fill_noalloc (the DP fill with its scratch in a static device array, no allocation inside the call) gives the
outputs of the existing entry point bt2gpu_test_fill, bit for bit (integer work: no tolerance), on two calls in a row with
different problems -- the second call finds the first one's contents in the shared scratch.  Called through the C-ABI with
device buffers that the TEST allocates (torch is only the allocator here).  Collected last on purpose (file name)."""
import ctypes as C
import numpy as np
import pytest
import common
from common import pkg

pytestmark = pytest.mark.gpu

MAXP = 128   # BT2GPU_NOALLOC_MAX_PROBLEMS (include/bt2gpu.h)


def _problems(rng, n, first):
    out = []
    for i in range(n):
        rows = first[i][0] if i < len(first) else int(rng.integers(1, 161))
        cols = first[i][1] if i < len(first) else int(rng.integers(1, 200))
        ref = rng.integers(0, 4, size=cols + 1).astype(np.uint8)
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
        out.append((rd.astype(np.uint8), qual, masks, rows, cols, -int(rng.integers(0, 120))))
    return out


def _pack(probs):
    n = len(probs)
    seq_off = np.zeros(n, np.uint32); rf_off = np.zeros(n, np.uint32); fl_off = np.zeros(n, np.uint64)
    so = ro = fo = 0
    for i, p in enumerate(probs):
        seq_off[i] = so; rf_off[i] = ro; fl_off[i] = fo
        so += p[3]; ro += p[4] + 1; fo += p[3] * p[4]
    return dict(n=n, seq_off=seq_off, rf_off=rf_off, fl_off=fl_off, nflags=fo,
                rows=np.array([p[3] for p in probs], np.uint16), cols=np.array([p[4] for p in probs], np.uint16),
                minsc=np.array([p[5] for p in probs], np.int32),
                seq=np.concatenate([p[0] for p in probs]), qual=np.concatenate([p[1] for p in probs]), rf=np.concatenate([p[2] for p in probs]))


def _existing(g, gpu, a):
    """bt2gpu_test_fill (mode 2: the production kernel in the lane layout the no-allocation entry point uses): host pointers in and out"""
    n, nf = a["n"], a["nflags"]
    H = np.zeros(nf, np.uint8); E = np.zeros(nf, np.uint8); F = np.zeros(nf, np.uint8); FL = np.zeros(nf, np.uint8)
    cand = np.zeros(256 * n, np.uint16); ncand = np.zeros(n, np.uint32); lrmax = np.zeros(n, np.int32); band = np.zeros(2 * n, np.int32)
    ptr = lambda x: x.ctypes.data_as(C.c_void_p)
    g.bt2gpu_test_fill.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 16 + [C.c_int32, C.c_void_p]
    rc = g.bt2gpu_test_fill(gpu.ptr, n, ptr(a["seq_off"]), ptr(a["seq"]), ptr(a["qual"]), ptr(a["rows"]), ptr(a["cols"]), ptr(a["rf_off"]),
                            ptr(a["rf"]), ptr(a["minsc"]), ptr(a["fl_off"]), ptr(H), ptr(E), ptr(F), ptr(FL), ptr(cand), ptr(ncand), ptr(lrmax), 2, ptr(band))
    assert rc == 0, g.bt2gpu_last_error(gpu.ptr)
    return FL, cand, ncand, lrmax


def _noalloc_prepare(a, torch, poison):
    n, nf = a["n"], a["nflags"]
    dev = torch.device("cuda", 0)
    up = lambda x: torch.from_numpy(x.view(np.uint8) if x.dtype != np.uint8 else x).to(dev)     # bytes of the array, as is
    d = {k: up(a[k]) for k in ("seq_off", "seq", "qual", "rows", "cols", "rf_off", "rf", "minsc", "fl_off")}
    o = {"fl": torch.full((nf,), poison, dtype=torch.uint8, device=dev), "cand": torch.zeros(256 * n * 2, dtype=torch.uint8, device=dev),
         "ncand": torch.full((n * 4,), poison, dtype=torch.uint8, device=dev), "lrmax": torch.full((n * 4,), poison, dtype=torch.uint8, device=dev)}
    return d, o


def _noalloc_launch(g, gpu, a, d, o, stream=None):
    """bt2gpu_fill_noalloc: device pointers in and out, work only ENQUEUED on `stream` (None: the legacy default stream)"""
    g.bt2gpu_fill_noalloc.restype = C.c_int
    g.bt2gpu_fill_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32] + [C.c_void_p] * 13
    rc = g.bt2gpu_fill_noalloc(gpu.ptr, C.c_void_p(stream.cuda_stream) if stream is not None else None, a["n"], d["seq_off"].data_ptr(), d["seq"].data_ptr(),
                               d["qual"].data_ptr(), d["rows"].data_ptr(), d["cols"].data_ptr(), d["rf_off"].data_ptr(), d["rf"].data_ptr(),
                               d["minsc"].data_ptr(), d["fl_off"].data_ptr(), o["fl"].data_ptr(), o["cand"].data_ptr(), o["ncand"].data_ptr(), o["lrmax"].data_ptr())
    assert rc == 0, g.bt2gpu_last_error(gpu.ptr)


def _collect(o):
    return (o["fl"].cpu().numpy(), o["cand"].cpu().numpy().view(np.uint16), o["ncand"].cpu().numpy().view(np.uint32), o["lrmax"].cpu().numpy().view(np.int32))


def _noalloc(g, gpu, a, torch, poison, stream=None):
    d, o = _noalloc_prepare(a, torch, poison)
    torch.cuda.synchronize()
    _noalloc_launch(g, gpu, a, d, o, stream)
    torch.cuda.synchronize()
    return _collect(o)


def _same(a, got, want, what):
    FL, cand, ncand, lrmax = want
    gFL, gcand, gncand, glrmax = got
    assert np.array_equal(glrmax, lrmax), what + ": lrmax"
    assert np.array_equal(gncand, ncand), what + ": ncand"
    for i in range(a["n"]):
        assert np.array_equal(gcand[256 * i:256 * i + ncand[i]], cand[256 * i:256 * i + ncand[i]]), "%s problem %d: candidates" % (what, i)
    assert np.array_equal(gFL, FL), what + ": flags"


def test_fill_noalloc_calls_on_two_streams_are_ordered():
    """Two calls enqueued back to back on two different non-default streams, nothing synchronised in between: they share the static scratch,
    so the second one's kernels must wait for the first one's (the library orders them with an event) -- both results are right."""
    torch = pytest.importorskip("torch")
    job = pkg.Job(common.index_base("lambda"), "")
    gpu = pkg.GpuAligner(job, 0)
    g = pkg.load_gpu()
    rng = np.random.default_rng(77)
    a1 = _pack(_problems(rng, MAXP, [(160, 199), (150, 199)])); a2 = _pack(_problems(rng, 90, [(150, 199), (33, 60)]))
    w1, w2 = _existing(g, gpu, a1), _existing(g, gpu, a2)
    s1, s2 = torch.cuda.Stream(device=0), torch.cuda.Stream(device=0)
    for rep in range(3):
        (d1, o1), (d2, o2) = _noalloc_prepare(a1, torch, 0x11), _noalloc_prepare(a2, torch, 0x22)
        torch.cuda.synchronize()
        _noalloc_launch(g, gpu, a1, d1, o1, s1)
        _noalloc_launch(g, gpu, a2, d2, o2, s2)        # enqueued while the first call's kernels are still running
        torch.cuda.synchronize()
        _same(a1, _collect(o1), w1, "stream 1 (rep %d)" % rep); _same(a2, _collect(o2), w2, "stream 2 (rep %d)" % rep)
    gpu.close()


def test_fill_noalloc_equals_existing_entry_point_twice():
    torch = pytest.importorskip("torch")
    job = pkg.Job(common.index_base("lambda"), "")
    gpu = pkg.GpuAligner(job, 0)
    g = pkg.load_gpu()
    rng = np.random.default_rng(2024)
    # first call: the full 128 problems, big ones first (fills the scratch); second call: fewer, smaller, other sizes
    # third call: on a stream of the caller's, after the two on the default stream
    calls = [_problems(rng, MAXP, [(160, 199), (150, 199), (129, 150), (128, 199), (65, 70), (64, 100), (1, 1), (2, 5)]),
             _problems(rng, 37, [(1, 3), (33, 40), (100, 120), (160, 1)]),
             _problems(rng, 61, [(150, 199), (7, 190)])]
    side = torch.cuda.Stream(device=0)
    for k, probs in enumerate(calls):
        a = _pack(probs)
        FL, cand, ncand, lrmax = _existing(g, gpu, a)
        gFL, gcand, gncand, glrmax = _noalloc(g, gpu, a, torch, poison=0xA5 if k == 0 else 0x5A, stream=side if k == 2 else None)
        assert np.array_equal(glrmax, lrmax), "call %d: lrmax" % k
        assert np.array_equal(gncand, ncand), "call %d: ncand" % k
        for i in range(a["n"]):
            assert np.array_equal(gcand[256 * i:256 * i + ncand[i]], cand[256 * i:256 * i + ncand[i]]), "call %d problem %d: candidates" % (k, i)
        if not np.array_equal(gFL, FL):
            j = int(np.nonzero(gFL != FL)[0][0]); i = int(np.searchsorted(a["fl_off"], j, side="right") - 1)
            raise AssertionError("call %d problem %d (%dx%d): flag byte %d differs: %02x vs %02x" %
                                 (k, i, a["rows"][i], a["cols"][i], j - int(a["fl_off"][i]), gFL[j], FL[j]))
    # a problem whose size is out of range (161 rows): switched off, its outputs say so and its flags are not touched
    probs = _problems(rng, 3, [(20, 30), (160, 199), (40, 60)])
    a = _pack(probs)
    a["rows"][1] = 161
    gFL, gcand, gncand, glrmax = _noalloc(g, gpu, a, torch, poison=0x77)
    assert gncand[1] == 0 and glrmax[1] == -1
    lo, hi = int(a["fl_off"][1]), int(a["fl_off"][2])
    assert (gFL[lo:hi] == 0x77).all() and (gFL[:lo] != 0x77).any()
    # beyond the static scratch: rejected with a message, nothing launched
    a = _pack(_problems(rng, 4, []))
    g.bt2gpu_last_error.restype = C.c_char_p
    rc = g.bt2gpu_fill_noalloc(gpu.ptr, None, MAXP + 1, *([C.c_void_p(8)] * 13))
    assert rc != 0 and b"at most 128" in g.bt2gpu_last_error(gpu.ptr)
    gpu.close()
