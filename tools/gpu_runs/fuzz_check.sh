#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_fuzz_options.py -q -m gpu 2>&1 | tail -12
