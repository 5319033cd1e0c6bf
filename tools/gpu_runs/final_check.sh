#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) > gpurun_out/c44_pytest.log 2>&1; tail -8 gpurun_out/c44_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py > gpurun_out/c44_bench.json 2> gpurun_out/c44_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c44_bench.json").read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["kernel"], round(d["roofline"]["frac"],3), round(d["cpu_baseline"]["value"]), d["clocks"])
print("   f2f", round(f["value"]), f["reader_sink_threads"], f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
