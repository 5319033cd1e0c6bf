#!/bin/bash
# SURVEY 8(d) configs beyond C2 on one GPU: parity sample vs the unmodified reference binary + throughput
mkdir -p gpurun_out
( time python tools/scale_check.py --ref-mbp 100 --reads 1000000 --sample 100000 --i16 --tag c3_100mbp_sensitive ) > gpurun_out/r2_c3.json 2> gpurun_out/r2_c3.err; echo "c3 rc=$?"
( time python tools/scale_check.py --ref-mbp 100 --reads 1000000 --sample 100000 --repeats 0.1 --opts "--very-sensitive --dpad 12" --tag c4rep_100mbp_verysensitive ) > gpurun_out/r2_c4rep100.json 2> gpurun_out/r2_c4rep100.err; echo "c4rep100 rc=$?"
tail -c 1500 gpurun_out/r2_c3.json; tail -3 gpurun_out/r2_c3.err; tail -c 1500 gpurun_out/r2_c4rep100.json; tail -3 gpurun_out/r2_c4rep100.err
