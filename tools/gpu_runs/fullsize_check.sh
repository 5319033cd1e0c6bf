#!/bin/bash
python -m pytest tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -3
