"""ctypes mirror of include/bt2gpu.h and include/bt2host.h."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_paths():
    """BT2GPU_LIB: another build of libbt2gpu.so to load instead (kernel experiments)."""
    return {"gpu": os.environ.get("BT2GPU_LIB") or os.path.join(_HERE, "libbt2gpu.so"), "host": os.path.join(_HERE, "libbt2host.so")}


def use_pinned_batches():
    """From now on the host library allocates the batches' base / quality arrays with libbt2gpu's bt2gpu_host_alloc (page-locked memory):
    bt2gpu_align_batch copies them to the device without a staging pass.  Needs a CUDA device; process-wide, once, before the first batch."""
    h, g = load_host(), load_gpu()
    g.bt2gpu_host_alloc.restype = C.c_void_p
    h.bt2host_set_array_allocator.argtypes = [C.c_void_p, C.c_void_p]
    return h.bt2host_set_array_allocator(C.cast(g.bt2gpu_host_alloc, C.c_void_p), C.cast(g.bt2gpu_host_free, C.c_void_p)) == 0


class bt2gpu_index(C.Structure):
    _fields_ = [("len", C.c_uint64), ("lineRate", C.c_int32), ("offRate", C.c_int32), ("ftabChars", C.c_int32),
                ("reserved0", C.c_int32), ("nPat", C.c_uint64), ("plen", C.c_void_p), ("nFrag", C.c_uint64),
                ("rstarts", C.c_void_p), ("ebwt", C.c_void_p), ("ebwtTotLen", C.c_uint64), ("zOff", C.c_uint64),
                ("fchr", C.c_void_p), ("ftab", C.c_void_p), ("ftabLen", C.c_uint64), ("eftab", C.c_void_p),
                ("eftabLen", C.c_uint64), ("offs", C.c_void_p), ("offsLen", C.c_uint64)]


class bt2gpu_ref(C.Structure):
    _fields_ = [("nrecs", C.c_uint64), ("rec_off", C.c_void_p), ("rec_len", C.c_void_p), ("rec_first", C.c_void_p),
                ("buf", C.c_void_p), ("bufBases", C.c_uint64)]


class bt2gpu_params(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("abi_version", "seedLen", "seedMms", "nSeedRounds", "seedBoostThresh", "maxhalf", "mxIter", "mxUg", "mxDp",
                 "streak", "doExtend", "enable8", "tighten", "khits", "mhits", "msample", "nofw", "norc", "gapbar",
                 "rdGapConst", "rdGapLinear", "rfGapConst", "rfGapLinear", "matchBonus", "i16", "exactBestLtMinsc")] + \
               [("reserved", C.c_int32 * 6), ("mmpens", C.c_int32 * 256), ("npens", C.c_int32 * 256)]


class bt2gpu_reads(C.Structure):
    _fields_ = [("n", C.c_uint32), ("seq_off", C.c_void_p), ("seq", C.c_void_p), ("qual", C.c_void_p),
                ("rng_seed", C.c_void_p), ("minsc", C.c_void_p), ("nceil", C.c_void_p), ("interval", C.c_void_p)]


class bt2gpu_edit(C.Structure):
    _fields_ = [("pos", C.c_uint32), ("chr", C.c_uint8), ("qchr", C.c_uint8), ("type", C.c_uint8), ("pad", C.c_uint8)]


class bt2gpu_aln(C.Structure):
    _fields_ = [("refoff", C.c_int64), ("tidx", C.c_uint32), ("score", C.c_int32), ("minsc", C.c_int32), ("ns", C.c_int32),
                ("gaps", C.c_int32), ("edits", C.c_int32), ("basesAligned", C.c_int32), ("refns", C.c_int32),
                ("edit_off", C.c_uint32), ("n_edits", C.c_uint16), ("fw", C.c_uint8), ("pad", C.c_uint8)]


class bt2gpu_read_result(C.Structure):
    _fields_ = [("aln_off", C.c_uint32), ("n_alns", C.c_uint32), ("rng_last", C.c_uint32), ("rng_lastOff", C.c_uint32),
                ("minsc", C.c_int32), ("status", C.c_uint8), ("flags", C.c_uint8), ("rounds", C.c_uint8), ("done_policy", C.c_uint8),
                ("nExIters", C.c_uint32), ("nExDps", C.c_uint32), ("nExDpSuccs", C.c_uint32), ("nExDpFails", C.c_uint32),
                ("nDpFail", C.c_uint32), ("nDpFailStreak", C.c_uint32), ("nDpLastSucc", C.c_uint32), ("nRedundants", C.c_uint32),
                ("bestLtMinsc", C.c_int32), ("nonzTot", C.c_uint32), ("numElts", C.c_uint32), ("numRanges", C.c_uint32),
                ("nSeedsTried", C.c_uint32), ("scratch_off", C.c_uint32), ("n_scratch", C.c_uint32),
                ("nExUgs", C.c_uint32), ("exhaustive", C.c_uint32)]


class bt2gpu_results(C.Structure):
    _fields_ = [("n", C.c_uint32), ("reads", C.POINTER(bt2gpu_read_result)), ("alns", C.POINTER(bt2gpu_aln)), ("n_alns", C.c_uint64),
                ("edits", C.POINTER(bt2gpu_edit)), ("n_edits", C.c_uint64)]


class bt2gpu_counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("reads", "seeds_searched", "lf_steps_search", "lf_steps_extend", "lf_sides_touched", "sa_rows_resolved",
                 "lf_steps_resolve", "dp_problems", "dp_cells", "bt_attempts", "bt_steps", "alns_reported")] + \
               [(n, C.c_double) for n in ("ms_seed", "ms_glue", "ms_resolve", "ms_fill", "ms_bt", "ms_h2d", "ms_d2h", "ms_run")] + \
               [(n, C.c_uint64) for n in ("launches", "n_seed", "n_step", "n_fill", "n_bt", "ext_text_steps", "dp_cells_computed", "sa_probes")]


READ_DT = np.dtype([(n, t) for n, t in
                    [("aln_off", "<u4"), ("n_alns", "<u4"), ("rng_last", "<u4"), ("rng_lastOff", "<u4"), ("minsc", "<i4"),
                     ("status", "u1"), ("flags", "u1"), ("rounds", "u1"), ("done_policy", "u1"),
                     ("nExIters", "<u4"), ("nExDps", "<u4"), ("nExDpSuccs", "<u4"), ("nExDpFails", "<u4"), ("nDpFail", "<u4"),
                     ("nDpFailStreak", "<u4"), ("nDpLastSucc", "<u4"), ("nRedundants", "<u4"), ("bestLtMinsc", "<i4"),
                     ("nonzTot", "<u4"), ("numElts", "<u4"), ("numRanges", "<u4"), ("nSeedsTried", "<u4"),
                     ("scratch_off", "<u4"), ("n_scratch", "<u4"), ("nExUgs", "<u4"), ("exhaustive", "<u4")]])
ALN_DT = np.dtype([("refoff", "<i8"), ("tidx", "<u4"), ("score", "<i4"), ("minsc", "<i4"), ("ns", "<i4"), ("gaps", "<i4"),
                   ("edits", "<i4"), ("basesAligned", "<i4"), ("refns", "<i4"), ("edit_off", "<u4"), ("n_edits", "<u2"),
                   ("fw", "u1"), ("pad", "u1")])
EDIT_DT = np.dtype([("pos", "<u4"), ("chr", "u1"), ("qchr", "u1"), ("type", "u1"), ("pad", "u1")])
assert READ_DT.itemsize == C.sizeof(bt2gpu_read_result) and ALN_DT.itemsize == C.sizeof(bt2gpu_aln) and EDIT_DT.itemsize == C.sizeof(bt2gpu_edit)

_host = None
_gpu = None
_gpu_long = None


def load_host():
    global _host
    if _host is None:
        h = C.CDLL(lib_paths()["host"])
        h.bt2host_job_create.restype = C.c_void_p
        h.bt2host_job_create.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
        h.bt2host_job_destroy.argtypes = [C.c_void_p]
        for f, rt in (("bt2host_job_index", C.POINTER(bt2gpu_index)), ("bt2host_job_ref", C.POINTER(bt2gpu_ref)),
                      ("bt2host_job_params", C.POINTER(bt2gpu_params))):
            getattr(h, f).restype = rt
            getattr(h, f).argtypes = [C.c_void_p]
        h.bt2host_batch_from_fastq.restype = C.c_void_p
        h.bt2host_batch_from_fastq.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64, C.c_char_p, C.c_size_t]
        h.bt2host_batch_name.restype = C.c_char_p
        h.bt2host_batch_name.argtypes = [C.c_void_p, C.c_uint32]
        h.bt2host_batch_from_fastq_range.restype = C.c_void_p
        h.bt2host_batch_from_fastq_range.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
        h.bt2host_batch_from_arrays.restype = C.c_void_p
        h.bt2host_batch_from_arrays.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_char_p, C.c_size_t]
        h.bt2host_batch_destroy.argtypes = [C.c_void_p]
        h.bt2host_batch_reads.restype = C.POINTER(bt2gpu_reads)
        h.bt2host_batch_reads.argtypes = [C.c_void_p]
        h.bt2host_sam_header.restype = C.c_void_p
        h.bt2host_sam_header.argtypes = [C.c_void_p]
        h.bt2host_sam_records.restype = C.c_void_p
        h.bt2host_sam_records.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(bt2gpu_results), C.c_int, C.POINTER(C.c_size_t)]
        h.bt2host_sam_write.restype = C.c_int
        h.bt2host_sam_write.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(bt2gpu_results), C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        h.bt2host_sam_format.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(bt2gpu_results), C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        h.bt2host_sam_flush.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        h.bt2host_free.argtypes = [C.c_void_p]
        h.bt2host_job_metrics.argtypes = [C.c_void_p, C.POINTER(C.c_uint64 * 4)]
        h.bt2host_job_set_threads.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _host = h
    return _host


def load_gpu(long_reads=False):
    """Load libbt2gpu.so -- or libbt2gpu_long.so, the same sources built with the reference's ALN_MAX_ROWS / ALN_MAX_COLS raised to
    1000 / 1072 (reads up to 1 kbp) -- ; fails loudly if it has not been built (there is no CPU path)."""
    global _gpu, _gpu_long
    if (_gpu_long if long_reads else _gpu) is None:
        p = lib_paths()["gpu"]
        if long_reads:
            p = p.replace("libbt2gpu.so", "libbt2gpu_long.so")
        if not os.path.exists(p):
            raise RuntimeError("%s is missing (run `make` or __graft_entry__.build()); bt2gpu has no CPU fallback" % os.path.basename(p))
        g = C.CDLL(p)
        g.bt2gpu_create.restype = C.c_void_p
        g.bt2gpu_create.argtypes = [C.POINTER(bt2gpu_index), C.POINTER(bt2gpu_ref), C.POINTER(bt2gpu_params), C.c_int, C.c_char_p, C.c_size_t]
        for f in ("bt2gpu_align_batch",):
            getattr(g, f).argtypes = [C.c_void_p, C.POINTER(bt2gpu_reads), C.POINTER(bt2gpu_results)]
        g.bt2gpu_submit.argtypes = [C.c_void_p, C.POINTER(bt2gpu_reads)]
        g.bt2gpu_wait.argtypes = [C.c_void_p, C.POINTER(bt2gpu_results)]
        g.bt2gpu_upload.argtypes = [C.c_void_p, C.POINTER(bt2gpu_reads)]
        g.bt2gpu_run_resident.argtypes = [C.c_void_p]
        g.bt2gpu_fetch.argtypes = [C.c_void_p, C.POINTER(bt2gpu_results)]
        g.bt2gpu_get_counters.argtypes = [C.c_void_p, C.POINTER(bt2gpu_counters)]
        g.bt2gpu_reset_counters.argtypes = [C.c_void_p]
        g.bt2gpu_last_error.restype = C.c_char_p
        g.bt2gpu_last_error.argtypes = [C.c_void_p]
        g.bt2gpu_destroy.argtypes = [C.c_void_p]
        g.bt2gpu_abi_version.restype = C.c_int
        if (g.bt2gpu_max_rows() > 160) != bool(long_reads):
            raise RuntimeError("%s was built for reads of up to %d bases" % (os.path.basename(p), g.bt2gpu_max_rows()))
        if long_reads:
            _gpu_long = g
        else:
            _gpu = g
    return _gpu_long if long_reads else _gpu


class Job:
    """A loaded .bt2l index + parsed bowtie2 options (libbt2host)."""

    def __init__(self, index_base, options=""):
        h = load_host()
        err = C.create_string_buffer(512)
        self._h = h
        self.ptr = h.bt2host_job_create(index_base.encode(), options.encode(), err, 512)
        if not self.ptr:
            raise RuntimeError("bt2host_job_create: " + err.value.decode())
        self.index = h.bt2host_job_index(self.ptr)
        self.ref = h.bt2host_job_ref(self.ptr)
        self.params = h.bt2host_job_params(self.ptr)

    def fastq(self, path, skip=0, max_reads=0):
        err = C.create_string_buffer(512)
        p = self._h.bt2host_batch_from_fastq(self.ptr, path.encode(), skip, max_reads, err, 512)
        if not p:
            raise RuntimeError("bt2host_batch_from_fastq: " + err.value.decode())
        return Batch(self, p)

    def fastq_range(self, path, offset, max_bytes):
        """The records lying completely inside [offset, offset + max_bytes) -> (batch, offset of the next record); n == 0 at the end."""
        err = C.create_string_buffer(512)
        nxt = C.c_uint64(0)
        p = self._h.bt2host_batch_from_fastq_range(self.ptr, path.encode(), offset, max_bytes, C.byref(nxt), err, 512)
        if not p:
            raise RuntimeError("bt2host_batch_from_fastq_range: " + err.value.decode())
        return Batch(self, p), nxt.value

    def arrays(self, seqs, quals=None, names=None):
        """Build a batch from python strings (ACGTN) and Phred+33 quality strings."""
        n = len(seqs)
        off = np.zeros(n + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(s) for s in seqs])
        sb = "".join(seqs).encode()
        qb = ("".join(quals) if quals is not None else "I" * len(sb)).encode()
        nm = None
        if names is not None:
            nm = (C.c_char_p * n)(*[x.encode() for x in names])
        err = C.create_string_buffer(512)
        p = self._h.bt2host_batch_from_arrays(self.ptr, n, off.ctypes.data, sb, qb, C.cast(nm, C.c_void_p) if nm is not None else None, err, 512)
        if not p:
            raise RuntimeError("bt2host_batch_from_arrays: " + err.value.decode())
        return Batch(self, p)

    def sam_header(self):
        p = self._h.bt2host_sam_header(self.ptr)
        s = C.string_at(p).decode()
        self._h.bt2host_free(p)
        return s

    def sam(self, batch, results, mode=1):
        n = C.c_size_t(0)
        p = self._h.bt2host_sam_records(self.ptr, batch.ptr, C.byref(results.raw), mode, C.byref(n))
        s = C.string_at(p, n.value).decode("latin-1")
        self._h.bt2host_free(p)
        return s

    def sam_bytes(self, batch, results, mode=1):
        """SAM records as bytes (no text decoding)."""
        n = C.c_size_t(0)
        p = self._h.bt2host_sam_records(self.ptr, batch.ptr, C.byref(results.raw), mode, C.byref(n))
        s = C.string_at(p, n.value)
        self._h.bt2host_free(p)
        return s

    def sam_write(self, batch, results, fileobj, mode=1):
        """SAM records written to an open binary file (bt2host_sam_write); returns the number of bytes."""
        fileobj.flush()
        n = C.c_size_t(0)
        if self._h.bt2host_sam_write(self.ptr, batch.ptr, C.byref(results.raw), mode, fileobj.fileno(), C.byref(n)) != 0:
            raise OSError("bt2host_sam_write failed")
        return n.value

    def sam_format(self, batch, results, block_set, mode=1):
        """First half of sam_write: the records formatted into the job's block set 0 or 1 (the batch and its results are free afterwards)."""
        n = C.c_size_t(0)
        if self._h.bt2host_sam_format(self.ptr, batch.ptr, C.byref(results.raw), mode, block_set, C.byref(n)) != 0:
            raise OSError("bt2host_sam_format failed")
        return n.value

    def sam_flush(self, block_set, fileobj):
        """Second half: block set 0 or 1 written to an open binary file; returns the number of bytes."""
        fileobj.flush()
        n = C.c_size_t(0)
        if self._h.bt2host_sam_flush(self.ptr, block_set, fileobj.fileno(), C.byref(n)) != 0:
            raise OSError("bt2host_sam_flush failed")
        return n.value

    def set_threads(self, reader=0, sink=0):
        """OpenMP threads of the FASTQ reader and of the SAM sink (0: all)."""
        self._h.bt2host_job_set_threads(self.ptr, reader, sink)

    def metrics(self):
        a = (C.c_uint64 * 4)()
        self._h.bt2host_job_metrics(self.ptr, C.byref(a))
        return list(a)

    def close(self):
        if self.ptr:
            self._h.bt2host_job_destroy(self.ptr)
            self.ptr = None


class Batch:
    def __init__(self, job, ptr):
        self.job, self.ptr = job, ptr
        self.reads = job._h.bt2host_batch_reads(ptr)

    @property
    def n(self):
        return self.reads.contents.n

    def name(self, i):
        p = self.job._h.bt2host_batch_name(self.ptr, i)
        return p.decode("latin-1") if p is not None else None

    def nbases(self):
        r = self.reads.contents
        off = np.ctypeslib.as_array(C.cast(r.seq_off, C.POINTER(C.c_uint32)), shape=(r.n + 1,))
        return int(off[-1])

    def close(self):
        if self.ptr:
            self.job._h.bt2host_batch_destroy(self.ptr)
            self.ptr = None


class Results:
    """Numpy views of a bt2gpu_results (valid until the producer's next call)."""

    def __init__(self, raw):
        self.raw = raw
        n = raw.n
        self.reads = np.ctypeslib.as_array(C.cast(raw.reads, C.POINTER(C.c_uint8)), shape=(n * READ_DT.itemsize,)).view(READ_DT) if n else np.zeros(0, READ_DT)
        self.alns = np.ctypeslib.as_array(C.cast(raw.alns, C.POINTER(C.c_uint8)), shape=(raw.n_alns * ALN_DT.itemsize,)).view(ALN_DT) if raw.n_alns else np.zeros(0, ALN_DT)
        self.edits = np.ctypeslib.as_array(C.cast(raw.edits, C.POINTER(C.c_uint8)), shape=(raw.n_edits * EDIT_DT.itemsize,)).view(EDIT_DT) if raw.n_edits else np.zeros(0, EDIT_DT)


class Counters(dict):
    pass


class GpuAligner:
    """bt2gpu context bound to one CUDA device (index and reference resident in HBM)."""

    def __init__(self, job, device=0, long_reads=False):
        g = load_gpu(long_reads)
        self._g = g
        err = C.create_string_buffer(512)
        self.ptr = g.bt2gpu_create(job.index, job.ref, job.params, device, err, 512)
        if not self.ptr:
            raise RuntimeError("bt2gpu_create: " + err.value.decode())
        self.job = job

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s: %s" % (what, self._g.bt2gpu_last_error(self.ptr).decode()))

    def align(self, batch):
        r = bt2gpu_results()
        self._ck(self._g.bt2gpu_align_batch(self.ptr, batch.reads, C.byref(r)), "bt2gpu_align_batch")
        return Results(r)

    def submit(self, batch):
        """bt2gpu_submit: returns at once, the batch is aligned on the context's worker thread; keep `batch` alive until wait()."""
        self._ck(self._g.bt2gpu_submit(self.ptr, batch.reads), "bt2gpu_submit")

    def wait(self):
        """bt2gpu_wait: the results of the submitted batch (valid until the next-but-one align / submit on this context)."""
        r = bt2gpu_results()
        self._ck(self._g.bt2gpu_wait(self.ptr, C.byref(r)), "bt2gpu_wait")
        return Results(r)

    def align_raw(self, reads_struct):
        """bt2gpu_align_batch on a caller-built bt2gpu_reads (e.g. views over pinned host memory)."""
        r = bt2gpu_results()
        self._ck(self._g.bt2gpu_align_batch(self.ptr, C.byref(reads_struct), C.byref(r)), "bt2gpu_align_batch")
        return Results(r)

    def upload(self, batch):
        self._ck(self._g.bt2gpu_upload(self.ptr, batch.reads), "bt2gpu_upload")

    def run_resident(self):
        self._ck(self._g.bt2gpu_run_resident(self.ptr), "bt2gpu_run_resident")

    def fetch(self):
        r = bt2gpu_results()
        self._ck(self._g.bt2gpu_fetch(self.ptr, C.byref(r)), "bt2gpu_fetch")
        return Results(r)

    def create_info(self):
        """bt2gpu_get_create_info: what bt2gpu_create spent (ms) and holds (bytes)"""
        ms = (C.c_double * 4)(); by = (C.c_uint64 * 4)()
        self._g.bt2gpu_get_create_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self._g.bt2gpu_get_create_info(self.ptr, ms, by)
        return {"ms": dict(zip(("upload", "suffix_array", "prefix_table", "slots"), [round(x, 1) for x in ms])),
                "bytes": dict(zip(("index_and_reference", "suffix_array", "prefix_table", "slots"), list(by)))}

    def counters(self, reset=False):
        c = bt2gpu_counters()
        self._g.bt2gpu_get_counters(self.ptr, C.byref(c))
        out = Counters((n, getattr(c, n)) for n, _ in bt2gpu_counters._fields_)
        if reset:
            self._g.bt2gpu_reset_counters(self.ptr)
        return out

    def close(self):
        if self.ptr:
            self._g.bt2gpu_destroy(self.ptr)
            self.ptr = None
