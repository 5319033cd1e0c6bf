"""bt2gpu: B200 (sm_100a) implementation of the seed-and-extend hot path of
sfiligoi/omp-bowtie2-prime behind a C-ABI (include/bt2gpu.h).

This package is a thin ctypes layer over two in-tree shared libraries:
  libbt2gpu.so   the CUDA kernels + C-ABI (the product)
  libbt2host.so  host-side .bt2l loader / FASTQ reader / SAM sink used to drive it
There is no CPU implementation of the path here: without the CUDA library and a
GPU, GpuAligner raises.
"""
from .binding import (Job, Batch, GpuAligner, Results, Counters, lib_paths, load_host, load_gpu, use_pinned_batches,
                      bt2gpu_reads, bt2gpu_results, bt2gpu_read_result, bt2gpu_aln, bt2gpu_edit,
                      bt2gpu_params, bt2gpu_index, bt2gpu_ref, bt2gpu_counters)

__all__ = ["Job", "Batch", "GpuAligner", "Results", "Counters", "lib_paths", "load_host", "load_gpu", "use_pinned_batches"]
