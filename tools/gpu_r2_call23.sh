#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fill.py tests/test_zz_no_alloc.py tests/test_gpu_parity.py -q -x 2>&1 | tail -3
python bench.py --no-cpu-baseline > gpurun_out/c23_bench.json 2> gpurun_out/c23_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c23_bench.json").read().strip().split("\n")[-1])
f=d.get("file_to_file") or {}
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
print(json.dumps(f))
PY
BT2HOST_TRACE=1 python tools/f2f_probe.py 1000000 4 > gpurun_out/c23_probe.log 2> gpurun_out/c23_probe.err
cat gpurun_out/c23_probe.log
grep "sam:" gpurun_out/c23_probe.err | awk '{n++; if(n%10==0) print}' | head -12
