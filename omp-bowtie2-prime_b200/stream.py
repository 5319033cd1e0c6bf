"""FASTQ file -> SAM file as a four-stage pipeline on one GPU: a reader thread parses byte ranges of the file
(bt2host_batch_from_fastq_range), the calling thread aligns them (bt2gpu_align_batch), a sink thread formats the records
(bt2host_sam_format) and a writer thread copies the formatted blocks into the file (bt2host_sam_flush).  The stages of neighbouring
batches overlap: the reference's own driver has the same stages (`PatternSourceReadAhead`, the batch loop, `AlnSinkSam::append`,
`OutputQueue::flush`; bt2_search.cpp:2414-2862, pat.h, aln_sink.cpp, outq.cpp).

One context does it: a context keeps two sets of result arrays and successive bt2gpu_align_batch calls take them in turn
(include/bt2gpu.h), so the sink can still be reading batch i while batch i+1 is aligned; batch i+2 waits for the sink.  The job keeps
two sets of formatted blocks in the same way: set i & 1 is written while batch i+1 is formatted into the other."""
import os
import queue
import stat
import threading
import time

from .binding import GpuAligner


def _cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def fastq_to_sam(job, fastq, out_path, device=0, chunk_bytes=64 << 20, mode=1, aligners=None, start=0, end=None, header=True, stats=None,
                 threads=None, preallocate=None):
    """Returns (reads, bytes written).  `aligners`: a list with the GpuAligner of `job` to reuse (else one is created and closed here;
    two objects are taken alternately, for aligners whose results last only until their next call).
    stats: a dict that receives the seconds each stage was busy ("parse", "align", "sam" = formatting, "write": they overlap) and the
    number of ranges.
    threads: (reader, sink) OpenMP threads, the sink's count is taken by the formatter and by the writer each; default: three eighths of
    the cores each (5 + 5 + 5 of 16: measured 12.7 M reads/s against 11.0 M with 5 + 9 + 9), two left to the aligner's own host threads
    when there are at least eight.
    preallocate: reserve the output file's blocks ahead of the sink (posix_fallocate of 1.5x the FASTQ bytes on a helper thread, cut back
    to what was written at the end): on tmpfs the sink otherwise spends most of its time waiting for the kernel to allocate and zero the
    pages it copies into.  Default: off (BT2HOST_STREAM_PREALLOC=1 turns it on) -- measured on tmpfs it cost more than it saved: the
    reservation holds the file's lock and a core while the pipeline wants both (8.1 vs 8.8 M reads/s).
    start / end: byte range of the file to take (both at record starts): a shard of a multi-GPU job (reads shard by batch,
    every device holds a replica of the index, the per-shard SAM files are concatenated in shard order); header=False for shards > 0."""
    if os.environ.get("BT2HOST_STREAM_CHUNK_MB"):
        chunk_bytes = int(os.environ["BT2HOST_STREAM_CHUNK_MB"]) << 20
    if threads is None and os.environ.get("BT2HOST_STREAM_THREADS"):      # "reader,sink"
        threads = tuple(int(x) for x in os.environ["BT2HOST_STREAM_THREADS"].split(","))
    if threads is None:
        nc = _cores()
        avail = max(2, nc - 2 if nc >= 8 else nc)
        threads = (max(1, avail * 3 // 8), max(1, avail * 3 // 8))    # reader, and formatter = writer: three teams share the cores
    job.set_threads(*threads)
    want_prealloc = preallocate if preallocate is not None else os.environ.get("BT2HOST_STREAM_PREALLOC", "0") != "0"
    own = aligners is None
    if own:
        aligners = [GpuAligner(job, device)]
    parsed, done = queue.Queue(maxsize=2), queue.Queue(maxsize=2)
    free = [threading.Semaphore(1), threading.Semaphore(1)]   # result set k may be written again once the sink is through with what it held
    errors = []
    busy = {"parse": 0.0, "align": 0.0, "sam": 0.0, "write": 0.0, "ranges": 0, "threads": list(threads)}

    def reader():
        try:
            off = start
            while not errors and (end is None or off < end):
                t0 = time.perf_counter()
                b, nxt = job.fastq_range(fastq, off, chunk_bytes if end is None else min(chunk_bytes, end - off))
                busy["parse"] += time.perf_counter() - t0
                if b.n == 0:
                    b.close()
                    break
                parsed.put(b)
                off = nxt
        except Exception as e:  # reported by the caller's thread
            errors.append(e)
        parsed.put(None)

    total = [0, 0]
    formatted = queue.Queue(maxsize=1)
    blocks_free = [threading.Semaphore(1), threading.Semaphore(1)]   # block set s may be formatted into again once the writer has written it

    def sink():
        """formats batch i into block set i & 1 (the reference's AlnSinkSam::append); the batch and its results are free after that"""
        i = 0
        try:
            while True:
                item = done.get()
                if item is None:
                    break
                b, res, k = item
                try:
                    if not errors:
                        blocks_free[i & 1].acquire()
                        t0 = time.perf_counter()
                        job.sam_format(b, res, i & 1, mode)
                        busy["sam"] += time.perf_counter() - t0
                        total[0] += b.n
                        formatted.put(i & 1)
                        i += 1
                finally:
                    b.close()
                    free[k].release()
        except Exception as e:
            errors.append(e)
            while True:                     # keep the producer from blocking
                item = done.get()
                if item is None:
                    break
                item[0].close(); free[item[2]].release()
        formatted.put(None)

    def writer():
        """writes the formatted block sets in order (OutputQueue::flush) while the next batch is being formatted"""
        try:
            with open(out_path, "w+b") as f:      # read + write: the blocks are copied into the file through a shared mapping
                pre = None
                if want_prealloc and stat.S_ISREG(os.fstat(f.fileno()).st_mode):
                    span = (end if end is not None else os.path.getsize(fastq)) - start
                    est = span + span // 2 + (1 << 20)

                    def reserve(fd=f.fileno(), est=est):
                        try:
                            os.posix_fallocate(fd, 0, est)     # (releases the GIL; pages the writer touches first are simply skipped)
                        except OSError:
                            pass
                    pre = threading.Thread(target=reserve); pre.start()
                if header:
                    f.write(job.sam_header().encode("latin-1"))
                while True:
                    s_ = formatted.get()
                    if s_ is None:
                        break
                    try:
                        if not errors:
                            t0 = time.perf_counter()
                            total[1] += job.sam_flush(s_, f)
                            busy["write"] += time.perf_counter() - t0
                    finally:
                        blocks_free[s_].release()
                if pre is not None:
                    pre.join()
                    f.flush()
                    f.truncate(f.tell())          # (the reservation was an estimate)
        except Exception as e:
            errors.append(e)
            while formatted.get() is not None:
                pass
            for sem in blocks_free:
                sem.release()

    tr, ts, tw = threading.Thread(target=reader), threading.Thread(target=sink), threading.Thread(target=writer)
    tr.start(); ts.start(); tw.start()
    i = 0
    try:
        while True:
            b = parsed.get()
            if b is None:
                break
            if errors:
                b.close()
                continue
            k = i & 1
            free[k].acquire()                  # (the other context's results may still be being written)
            t0 = time.perf_counter()
            res = aligners[k % len(aligners)].align(b)
            busy["align"] += time.perf_counter() - t0
            busy["ranges"] += 1
            done.put((b, res, k))
            i += 1
    except Exception as e:
        errors.append(e)
        while parsed.get() is not None:
            pass
    done.put(None)
    tr.join(); ts.join(); tw.join()
    if own:
        for g in aligners:
            g.close()
    job.set_threads(0, 0)
    if errors:
        raise errors[0]
    if stats is not None:
        stats.update(busy)
    return total[0], total[1]
