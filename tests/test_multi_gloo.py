"""CPU, world_size 2 over gloo: the sharding driver (omp-bowtie2-prime_b200/shard.py) -- contiguous
read shards, metric all-reduce, rank-ordered SAM gather -- gives the single-process answer.
The aligner plugged in here is the oracle (test infrastructure); on GPUs it is GpuAligner.align."""
import os
import subprocess
import sys
import pytest
import common

WORKER = r'''
import os, sys, importlib
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
import common
shard = importlib.import_module("omp-bowtie2-prime_b200.shard")
dist.init_process_group("gloo")
job = common.pkg.Job(common.index_base("syn"), "-k 2")
orc = common.Oracle(job)
fq = common.reads_path("syn100")
sam, met = shard.align_sharded(job, fq, lambda b: orc.align(b, 1), 600, mode=1)
if dist.get_rank() == 0:
    open(sys.argv[2], "w", encoding="latin-1").write(sam)
    open(sys.argv[2] + ".met", "w").write(" ".join(map(str, met)))
dist.destroy_process_group()
'''


def test_shard_bounds():
    import importlib
    shard = importlib.import_module("omp-bowtie2-prime_b200.shard")
    for n in (0, 1, 7, 600, 1000001):
        for w in (1, 2, 3, 8):
            b = [shard.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


def test_two_ranks_equal_one(tmp_path):
    if not common.have_builder():
        pytest.skip("no builder")
    common.index_base("syn"); common.reads_path("syn100")  # build fixtures before the ranks race for them
    w = tmp_path / "worker.py"; w.write_text(WORKER)
    out2 = str(tmp_path / "two.sam")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                    "--master-port", "29533", str(w), common.ROOT, out2], check=True, env=env, timeout=300)
    job = common.pkg.Job(common.index_base("syn"), "-k 2")
    batch = job.fastq(common.reads_path("syn100"), 0, 600)
    orc = common.Oracle(job)
    one = job.sam(batch, orc.align(batch, 1), 1)
    assert open(out2, encoding="latin-1").read() == one
    assert [int(x) for x in open(out2 + ".met").read().split()] == job.metrics()
