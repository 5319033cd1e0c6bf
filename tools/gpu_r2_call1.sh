#!/bin/bash
# round-2 GPU call 1: sanity of the round-1 state + baselines + sanitizer evidence
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/c1_smi.txt 2>&1
nproc > gpurun_out/c1_nproc.txt; lscpu | head -20 >> gpurun_out/c1_nproc.txt
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c1_pytest.log 2>&1
( time timeout 600 python bench.py --steps 5 --warmup 3 ) > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/c1_bench_ref.json 2> gpurun_out/c1_bench_ref.err
( time timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_smoke.py 150 ) > gpurun_out/c1_memcheck.log 2>&1
( time timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_smoke.py 60 ) > gpurun_out/c1_racecheck.log 2>&1
( time timeout 600 compute-sanitizer --tool synccheck --print-limit 20 python tools/sanitize_smoke.py 60 ) > gpurun_out/c1_synccheck.log 2>&1
tail -3 gpurun_out/c1_pytest.log; tail -c 600 gpurun_out/c1_bench.json; tail -5 gpurun_out/c1_memcheck.log gpurun_out/c1_racecheck.log gpurun_out/c1_synccheck.log
