#!/bin/bash
mkdir -p gpurun_out
BT2HOST_TRACE=1 python tools/f2f_probe.py 1000000 2 > gpurun_out/c25_probe.log 2> gpurun_out/c25_probe.err
cat gpurun_out/c25_probe.log
grep "sam:" gpurun_out/c25_probe.err | tail -16
