#!/bin/bash
# launch list of the bench command (after it has exited 0 without ncu)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-file-to-file"
$CMD > gpurun_out/c37_plain.json 2> gpurun_out/c37_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_bench_steps2_warmup1.csv $CMD > gpurun_out/c37_ncu.log 2>&1
echo "rc=$?"; wc -l gpurun_out/r2_launches_bench_steps2_warmup1.csv
