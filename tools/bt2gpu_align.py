#!/usr/bin/env python
"""FASTQ -> SAM on one or more B200s (one process per GPU):
  python tools/bt2gpu_align.py -x INDEX -U reads.fq -S out.sam [bowtie2 options...]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bt2gpu_align.py ...
"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    a = sys.argv[1:]
    idx = a[a.index("-x") + 1]; fq = a[a.index("-U") + 1]; out = a[a.index("-S") + 1]
    for k in ("-x", "-U", "-S"):
        i = a.index(k); del a[i:i + 2]
    import torch, torch.distributed as dist
    pkg = importlib.import_module("omp-bowtie2-prime_b200")
    shard = importlib.import_module("omp-bowtie2-prime_b200.shard")
    world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    job = pkg.Job(idx, " ".join(a))
    if world == 1:   # one GPU: reader, aligner and sink run as a pipeline over byte ranges of the file
        stream = importlib.import_module("omp-bowtie2-prime_b200.stream")
        try:
            stream.fastq_to_sam(job, fq, out, device=local)
        except RuntimeError as e:
            if "serial parser" not in str(e):
                raise
            gpu = pkg.GpuAligner(job, local)   # CR line ends / blank lines: whole file through the serial parser
            batch = job.fastq(fq)
            res = gpu.align(batch)
            with open(out, "wb") as f:
                f.write(job.sam_header().encode("latin-1"))
                job.sam_write(batch, res, f, 1)
            batch.close(); gpu.close()
        met = job.metrics()
        print("%d reads; %d (%.2f%%) aligned 0 times; %d aligned exactly 1 time; %d aligned >1 times" %
              (met[0], met[1], 100.0 * met[1] / max(met[0], 1), met[2], met[3]), file=sys.stderr)
        return
    gpu = pkg.GpuAligner(job, local)
    n = sum(1 for _ in open(fq, "rb")) // 4
    sam, met = shard.align_sharded(job, fq, gpu.align, n, mode=1, device=torch.device("cuda", local))
    if sam is not None:
        with open(out, "w", encoding="latin-1") as f:
            f.write(job.sam_header()); f.write(sam)
        print("%d reads; %d (%.2f%%) aligned 0 times; %d aligned exactly 1 time; %d aligned >1 times" %
              (met[0], met[1], 100.0 * met[1] / max(met[0], 1), met[2], met[3]), file=sys.stderr)
    gpu.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
