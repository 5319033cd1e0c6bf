"""Multi-GPU driver: reads shard across ranks (one process per GPU, index replicated), no
collective on the data path.  The only communication is the end-of-run sum of
the ReportingMetrics counters (aln_sink.h:44-233; bt2_search.cpp:2867-2869 merges them across
threads) and, when one SAM is wanted, a gather of the per-rank record blocks to rank 0."""
import torch
import torch.distributed as dist


def shard_bounds(n, world, rank):
    """Contiguous shard [lo, hi) of n reads for `rank`; sizes differ by at most one."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_metrics(metrics, device=None):
    """Sum [reads, unaligned, aligned once, aligned >1] over ranks (NCCL on GPUs, gloo on CPU)."""
    t = torch.tensor([int(x) for x in metrics], dtype=torch.int64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [int(x) for x in t.tolist()]


def gather_sam(text, dst=0):
    """Concatenate the ranks' SAM record blocks on `dst` in rank order (= read order, shards are contiguous)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return text
    out = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(text, out, dst=dst)
    return "".join(out) if out is not None else None


def align_sharded(job, fastq, align_fn, total_reads, mode=1, device=None):
    """Each rank parses and aligns its own shard with `align_fn(batch) -> Results`.
    Returns (sam text on rank 0 else None, summed metrics)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    lo, hi = shard_bounds(total_reads, world, rank)
    batch = job.fastq(fastq, lo, hi - lo) if hi > lo else None
    text = ""
    if batch is not None and batch.n > 0:
        res = align_fn(batch)
        text = job.sam(batch, res, mode)
    metrics = allreduce_metrics(job.metrics(), device)
    sam = gather_sam(text)
    if batch is not None:
        batch.close()
    return sam, metrics
