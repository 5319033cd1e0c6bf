#!/bin/bash
# C4 (3.1 Gbp, --very-sensitive) with the end-of-round kernels: parity sample, throughput, create-time breakdown
mkdir -p gpurun_out
python tools/scale_check.py --ref-mbp 3100 --reads 1000000 --sample 50000 --seed 12347 --opts "--very-sensitive --dpad 12" --tag c4_3100mbp_verysensitive > gpurun_out/c41_c4.json 2> gpurun_out/c41_c4.err; echo "c4 rc=$?"; cat gpurun_out/c41_c4.json; tail -3 gpurun_out/c41_c4.err
