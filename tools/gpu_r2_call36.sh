#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) > gpurun_out/c36_pytest.log 2>&1; tail -8 gpurun_out/c36_pytest.log
python bench.py --no-cpu-baseline --no-file-to-file > gpurun_out/c36_bench.json 2> gpurun_out/c36_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c36_bench.json").read().strip().split("\n")[-1])
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["kernel"], round(d["roofline"]["frac"],3))
PY
python tools/scale_check.py --ref-mbp 100 --reads 10000000 --sample 100000 --i16 --tag c3_100mbp_10Mreads > gpurun_out/c36_c3_10m.json 2> gpurun_out/c36_c3_10m.err; echo "c3 rc=$?"; cat gpurun_out/c36_c3_10m.json; tail -2 gpurun_out/c36_c3_10m.err
