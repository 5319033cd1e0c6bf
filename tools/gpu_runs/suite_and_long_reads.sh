#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -6 ) > gpurun_out/c29_pytest.log 2>&1; tail -9 gpurun_out/c29_pytest.log
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 8000 --long 170,200,230,255 --opts=--sensitive --tag c5_long255_100mbp --keep > gpurun_out/c29_c5_255.json 2> gpurun_out/c29_c5_255.err; echo "c5-255 rc=$?"; cat gpurun_out/c29_c5_255.json
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 4000 --long 250,500,750,1000 --opts=--sensitive --tag c5_long_100mbp > gpurun_out/c29_c5.json 2> gpurun_out/c29_c5.err; echo "c5 rc=$?"; cat gpurun_out/c29_c5.json
