"""CPU: the C-ABI libraries load and export every symbol the headers declare."""
import ctypes
import os
import re
import common
from common import pkg


def _declared(header):
    txt = open(os.path.join(common.ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(bt2(?:gpu|host)_[a-z0-9_]+)\s*\(", txt)))


def test_bt2gpu_exports():
    p = pkg.lib_paths()["gpu"]
    assert os.path.exists(p), "libbt2gpu.so not built (run make / __graft_entry__.build())"
    lib = ctypes.CDLL(p)
    names = _declared("bt2gpu.h")
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), n
    lib.bt2gpu_abi_version.restype = ctypes.c_int
    assert lib.bt2gpu_abi_version() == 4  # BT2GPU_ABI_VERSION in include/bt2gpu.h


def test_bt2host_exports():
    lib = ctypes.CDLL(pkg.lib_paths()["host"])
    for n in _declared("bt2host.h"):
        assert hasattr(lib, n), n


def test_struct_sizes_match_header():
    # the ctypes mirrors must have the C layout (checked against sizes compiled from the header)
    import subprocess, tempfile, textwrap
    src = textwrap.dedent('''
        #include <stdio.h>
        #include "bt2gpu.h"
        int main(){ printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(bt2gpu_index), sizeof(bt2gpu_ref), sizeof(bt2gpu_params),
          sizeof(bt2gpu_reads), sizeof(bt2gpu_edit), sizeof(bt2gpu_aln), sizeof(bt2gpu_read_result), sizeof(bt2gpu_counters)); return 0; }''')
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.run(["gcc", "-I", os.path.join(common.ROOT, "include"), "-o", os.path.join(d, "s"), os.path.join(d, "s.c")], check=True)
        out = subprocess.run([os.path.join(d, "s")], capture_output=True, text=True, check=True).stdout.split()
    got = [ctypes.sizeof(x) for x in (pkg.bt2gpu_index, pkg.bt2gpu_ref, pkg.bt2gpu_params, pkg.bt2gpu_reads, pkg.bt2gpu_edit,
                                      pkg.bt2gpu_aln, pkg.bt2gpu_read_result, pkg.bt2gpu_counters)]
    assert got == [int(x) for x in out]


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return
    if not common.have_builder():
        return
    job = pkg.Job(common.index_base("lambda"), "")
    try:
        pkg.GpuAligner(job)
    except RuntimeError as e:
        assert "no CUDA device" in str(e)
    else:
        raise AssertionError("GpuAligner must not work without a GPU")


def test_fastq_fast_path_equals_serial_parser(tmp_path):
    """the OpenMP whole-file FASTQ reader and the serial one (taken for CR line ends) build the same batch"""
    import numpy as np
    import common
    from common import pkg
    if not common.have_builder():
        import pytest
        pytest.skip("no builder")
    src = common.reads_path("lambda_reads_1k")
    crlf = tmp_path / "crlf.fq"
    crlf.write_bytes(open(src, "rb").read().replace(b"\n", b"\r\n"))
    for opts in ("", "--trim5 3 --trim3 7"):
        job = pkg.Job(common.index_base("lambda"), opts)
        a, b = job.fastq(src), job.fastq(str(crlf))
        ra, rb = a.reads.contents, b.reads.contents
        assert ra.n == rb.n == 1000
        for f, dt, cnt in (("seq_off", np.uint32, ra.n + 1), ("rng_seed", np.uint32, ra.n), ("minsc", np.int32, ra.n), ("interval", np.int32, ra.n)):
            xa = np.ctypeslib.as_array(ctypes.cast(getattr(ra, f), ctypes.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(cnt,))
            xb = np.ctypeslib.as_array(ctypes.cast(getattr(rb, f), ctypes.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(cnt,))
            assert (xa == xb).all(), f
        nb = a.nbases()
        for f in ("seq", "qual"):
            xa = np.ctypeslib.as_array(ctypes.cast(getattr(ra, f), ctypes.POINTER(ctypes.c_uint8)), shape=(nb,))
            xb = np.ctypeslib.as_array(ctypes.cast(getattr(rb, f), ctypes.POINTER(ctypes.c_uint8)), shape=(nb,))
            assert (xa == xb).all(), f
        # part of the file: skip / max_reads
        c, d = job.fastq(src, 100, 250), job.fastq(str(crlf), 100, 250)
        assert c.reads.contents.n == d.reads.contents.n == 250 and c.nbases() == d.nbases()


def test_fastq_fast_path_odd_records(tmp_path):
    """records off the usual shape: lower case, '.', characters the reference drops from the sequence line (pat.cpp:1151-1163), a read without
    a name, a name with '/', a quality line shorter than the read, --phred64; fast reader == serial reader, names and seeds included"""
    import random
    import numpy as np
    import common
    from common import pkg
    if not common.have_builder():
        import pytest
        pytest.skip("no builder")
    rnd = random.Random(5)
    recs = []
    for i in range(600):
        L = rnd.randint(20, 170)
        s = "".join(rnd.choice("ACGT") for _ in range(L))
        q = "".join(chr(rnd.randint(64, 104)) for _ in range(L))
        k = i % 8
        if k == 1:
            s = s.lower()
        elif k == 2:
            s = s[:5] + "." + s[6:11] + "n" + s[12:]
        elif k == 3:
            s = s[:7] + "-*1" + s[7:]          # dropped by the reader: the read keeps L bases, qualities index the kept ones
            q = q + "hhh"
        elif k == 4:
            q = q[:L - 4]
        name = "" if k == 5 else ("frag%d/1 extra" % i if k == 6 else "r%d" % i)
        recs.append("@%s\n%s\n+\n%s\n" % (name, s, q))
    lf = tmp_path / "odd.fq"; lf.write_text("".join(recs))
    crlf = tmp_path / "odd_crlf.fq"; crlf.write_bytes(lf.read_bytes().replace(b"\n", b"\r\n"))
    for opts in ("--phred64", "--phred64 --trim5 4 --trim3 9", "--phred64 --seed 7"):
        job = pkg.Job(common.index_base("lambda"), opts)
        a, b = job.fastq(str(lf)), job.fastq(str(crlf))
        ra, rb = a.reads.contents, b.reads.contents
        assert ra.n == rb.n == 600
        for f, dt, cnt in (("seq_off", np.uint32, ra.n + 1), ("rng_seed", np.uint32, ra.n), ("minsc", np.int32, ra.n), ("nceil", np.int32, ra.n)):
            xa = np.ctypeslib.as_array(ctypes.cast(getattr(ra, f), ctypes.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(cnt,))
            xb = np.ctypeslib.as_array(ctypes.cast(getattr(rb, f), ctypes.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(cnt,))
            assert (xa == xb).all(), (opts, f, int((xa != xb).argmax()))
        nb = a.nbases()
        assert nb == b.nbases()
        for f in ("seq", "qual"):
            xa = np.ctypeslib.as_array(ctypes.cast(getattr(ra, f), ctypes.POINTER(ctypes.c_uint8)), shape=(nb,))
            xb = np.ctypeslib.as_array(ctypes.cast(getattr(rb, f), ctypes.POINTER(ctypes.c_uint8)), shape=(nb,))
            assert (xa == xb).all(), (opts, f)
        assert [a.name(i) for i in range(600)] == [b.name(i) for i in range(600)]
        assert a.name(5) == "5" and a.name(6).startswith("frag6/1")
        a.close(); b.close(); job.close()


def test_fastq_ranges_equal_whole_file():
    """bt2host_batch_from_fastq_range: the batches of consecutive byte ranges, concatenated, are the whole-file batch"""
    import numpy as np
    import pytest
    if not common.have_builder():
        pytest.skip("no builder")
    src = common.reads_path("lambda_reads_1k")
    job = pkg.Job(common.index_base("lambda"), "--trim5 2")
    whole = job.fastq(src)
    rw = whole.reads.contents

    def arr(r, f, dt, cnt):
        return np.ctypeslib.as_array(ctypes.cast(getattr(r, f), ctypes.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(cnt,)).copy()
    for chunk in (3000, 50000, 10 ** 9):
        off, n, seq, qual, seed, names, lens = 0, 0, [], [], [], [], []
        while True:
            b, nxt = job.fastq_range(src, off, chunk)
            r = b.reads.contents
            if r.n == 0:
                assert nxt == off
                break
            so = arr(r, "seq_off", np.uint32, r.n + 1)
            lens.append(np.diff(so)); seq.append(arr(r, "seq", np.uint8, b.nbases())); qual.append(arr(r, "qual", np.uint8, b.nbases()))
            seed.append(arr(r, "rng_seed", np.uint32, r.n)); names += [b.name(i) for i in range(r.n)]
            n += r.n; assert nxt > off; off = nxt; b.close()
        assert n == rw.n and off == __import__("os").path.getsize(src)
        assert (np.concatenate(lens) == np.diff(arr(rw, "seq_off", np.uint32, rw.n + 1))).all()
        assert (np.concatenate(seq) == arr(rw, "seq", np.uint8, whole.nbases())).all()
        assert (np.concatenate(qual) == arr(rw, "qual", np.uint8, whole.nbases())).all()
        assert (np.concatenate(seed) == arr(rw, "rng_seed", np.uint32, rw.n)).all()
        assert names == [whole.name(i) for i in range(rw.n)]
    with pytest.raises(RuntimeError, match="longer than the read range"):
        job.fastq_range(src, 0, 20)


def test_array_allocator_hook_and_thread_split():
    """bt2host_set_array_allocator: the batches' base / quality arrays come from the caller's pair of functions (libbt2gpu's page-locked
    allocator in production; libc's malloc / free here), once per process; bt2host_job_set_threads is accepted and changes no result"""
    import common
    from common import pkg
    if not common.have_builder():
        import pytest
        pytest.skip("no builder")
    h = pkg.load_host()
    libc = ctypes.CDLL(None)
    h.bt2host_set_array_allocator.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    job = pkg.Job(common.index_base("lambda"), "")
    before = job.fastq(common.reads_path("lambda_reads_1k"))
    rc = h.bt2host_set_array_allocator(ctypes.cast(libc.malloc, ctypes.c_void_p), ctypes.cast(libc.free, ctypes.c_void_p))
    assert rc in (0, -1)                                   # (-1: another test of this process registered one already)
    assert h.bt2host_set_array_allocator(ctypes.cast(libc.malloc, ctypes.c_void_p), ctypes.cast(libc.free, ctypes.c_void_p)) == -1
    assert h.bt2host_set_array_allocator(None, None) == -1
    job.set_threads(2, 3)
    after = job.fastq(common.reads_path("lambda_reads_1k"))
    n = before.nbases()
    a = ctypes.string_at(before.reads.contents.seq, n); b = ctypes.string_at(after.reads.contents.seq, n)
    assert a == b and after.n == before.n == 1000
    job.set_threads(0, 0)
    after.close(); job.close()     # (`before` was allocated by malloc before the hook changed hands and is freed by free: same functions)
    before.close()
