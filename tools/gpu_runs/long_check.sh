#!/bin/bash
python -m pytest tests/test_long_reads.py tests/test_fuzz_options.py -q -m gpu -k long 2>&1 | tail -2
