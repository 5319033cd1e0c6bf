#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) > gpurun_out/c40_pytest.log 2>&1; tail -8 gpurun_out/c40_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --impl reference > gpurun_out/c40_bench_ref.json 2> gpurun_out/c40_bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/c40_bench.json 2> gpurun_out/c40_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c40_bench.json").read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["kernel"], round(d["roofline"]["frac"],3), round(d["cpu_baseline"]["value"]), d["clocks"])
print("   f2f", round(f["value"]), f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
r=json.loads(open("gpurun_out/c40_bench_ref.json").read().strip().split("\n")[-1])
print("reference arm", round(r["value"]), r["cpu_baseline"].get("wall_reads_per_s"), r["cpu_baseline"].get("threads"), r["cpu_baseline"].get("p"))
PY
