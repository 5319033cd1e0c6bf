#!/bin/bash
# ncu --set full of K_fill on a short bench run (the same command has just exited 0 without ncu, && below)
mkdir -p gpurun_out
CMD="python bench.py --reads 300000 --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fill -s 0 -c 1 -f -o gpurun_out/r2_fill $CMD > gpurun_out/ncu_fill.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu_fill.log; ls -la gpurun_out/
