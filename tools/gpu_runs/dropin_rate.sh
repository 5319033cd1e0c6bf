#!/bin/bash
mkdir -p gpurun_out
python tools/dropin_rate.py 400000 > gpurun_out/dropin_rate.json 2> gpurun_out/dropin_rate.err; echo rc=$?; cat gpurun_out/dropin_rate.json; tail -3 gpurun_out/dropin_rate.err
