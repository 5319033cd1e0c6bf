#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_dropin.py -x -q ) > gpurun_out/c6_dropin.log 2>&1; echo "dropin rc=$?"
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?"
( time timeout 900 python bench.py --steps 5 --warmup 3 ) > gpurun_out/c6_bench.json 2> gpurun_out/c6_bench.err; echo "bench rc=$?"
tail -15 gpurun_out/c6_dropin.log; tail -8 gpurun_out/c6_pytest.log; tail -c 2500 gpurun_out/c6_bench.json; tail -5 gpurun_out/c6_bench.err
