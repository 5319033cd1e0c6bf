#!/bin/bash
mkdir -p gpurun_out
lscpu | grep -E "Model name|Thread|Core|Socket|NUMA node\(s\)|^CPU\(s\)" 
BT2HOST_TRACE=1 python tools/f2f_probe.py 1000000 4 > gpurun_out/c22_probe.log 2> gpurun_out/c22_probe.err
cat gpurun_out/c22_probe.log
grep "sam:" gpurun_out/c22_probe.err | awk '{n++; if(n%8==0) print}' | head -30
grep "fastq:" gpurun_out/c22_probe.err | awk '{n++; if(n%8==0) print}' | head -30
