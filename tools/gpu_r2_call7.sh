#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?"
( time timeout 900 python bench.py --steps 5 --warmup 3 ) > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err; echo "bench rc=$?"
tail -30 gpurun_out/c7_pytest.log | cut -c1-300; tail -c 600 gpurun_out/c7_bench.json; tail -3 gpurun_out/c7_bench.err
