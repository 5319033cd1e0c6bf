"""FASTQ file -> SAM file as a three-stage pipeline on one GPU: a reader thread parses byte ranges of the file
(bt2host_batch_from_fastq_range), the calling thread aligns them (bt2gpu_align_batch), a sink thread formats and writes the
records (bt2host_sam_write).  The stages of neighbouring batches overlap: the reference's own driver has the same three
stages (`PatternSourceReadAhead`, the batch loop, `OutputQueue`; bt2_search.cpp:2414-2862, pat.h, outq.cpp).

Two contexts on the same device take the batches alternately, because a context's results stay valid only until its next
call and the sink is still reading batch i while batch i+1 is aligned."""
import queue
import threading
import time

from .binding import GpuAligner


def fastq_to_sam(job, fastq, out_path, device=0, chunk_bytes=64 << 20, mode=1, aligners=None, start=0, end=None, header=True, stats=None):
    """Returns (reads, bytes written).  `aligners`: two GpuAligner of `job` to reuse (else created and closed here).
    stats: a dict that receives the seconds each stage was busy ("parse", "align", "sam": they overlap) and the number of ranges.
    start / end: byte range of the file to take (both at record starts): a shard of a multi-GPU job (reads shard by batch,
    every device holds a replica of the index, the per-shard SAM files are concatenated in shard order); header=False for shards > 0."""
    own = aligners is None
    if own:
        aligners = [GpuAligner(job, device), GpuAligner(job, device)]
    parsed, done = queue.Queue(maxsize=2), queue.Queue(maxsize=2)
    free = [threading.Semaphore(1), threading.Semaphore(1)]   # context k may align again once the sink has written its last results
    errors = []
    busy = {"parse": 0.0, "align": 0.0, "sam": 0.0, "ranges": 0}

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

    def sink():
        try:
            with open(out_path, "wb") as f:
                if header:
                    f.write(job.sam_header().encode("latin-1"))
                while True:
                    item = done.get()
                    if item is None:
                        break
                    b, res, k = item
                    try:
                        if not errors:
                            t0 = time.perf_counter()
                            total[1] += job.sam_write(b, res, f, mode)
                            busy["sam"] += time.perf_counter() - t0
                            total[0] += b.n
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

    tr, ts = threading.Thread(target=reader), threading.Thread(target=sink)
    tr.start(); ts.start()
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
            res = aligners[k].align(b)
            busy["align"] += time.perf_counter() - t0
            busy["ranges"] += 1
            done.put((b, res, k))
            i += 1
    except Exception as e:
        errors.append(e)
        while parsed.get() is not None:
            pass
    done.put(None)
    tr.join(); ts.join()
    if own:
        for g in aligners:
            g.close()
    if errors:
        raise errors[0]
    if stats is not None:
        stats.update(busy)
    return total[0], total[1]
