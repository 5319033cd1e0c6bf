#!/bin/bash
mkdir -p gpurun_out
python bench.py --config c3 --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo rc=$?
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c3.json").read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(d["config"]["workload"][:80], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
if f: print("   f2f", round(f["value"]), "devnull", round(f["sam_to_dev_null"]["value"]))
PY
tail -2 gpurun_out/bench_c3.err
