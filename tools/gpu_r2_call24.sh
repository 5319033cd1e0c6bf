#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_long_reads.py tests/test_gpu_fill.py -q -x -m gpu 2>&1 | tail -15
python tests/host_rate.py 1000000 /dev/shm/hr --gpu 2>&1 | grep -E "threads" | tail -12
rm -rf /dev/shm/hr
