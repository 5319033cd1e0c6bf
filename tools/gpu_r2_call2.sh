#!/bin/bash
# round-2 GPU call 2: banded diagonal-sweep K_fill -- parity first, then numbers
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_gpu_fill.py tests/test_zz_no_alloc.py -x -q -s ) > gpurun_out/c2_fill.log 2>&1
echo "fill tests rc=$?" >> gpurun_out/c2_fill.log
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c2_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
( time timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ) > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err
tail -5 gpurun_out/c2_fill.log; tail -5 gpurun_out/c2_pytest.log; tail -c 1500 gpurun_out/c2_bench.json; tail -3 gpurun_out/c2_bench.err
