#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -5 ) > gpurun_out/c33_pytest.log 2>&1; tail -8 gpurun_out/c33_pytest.log
python bench.py > gpurun_out/c33_bench.json 2> gpurun_out/c33_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c33_bench.json").read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["kernel"], round(d["roofline"]["frac"],3), d["cpu_baseline"]["value"])
print("   f2f", round(f["value"]), f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
CMD="python bench.py --reads 300000 --steps 1 --warmup 1 --no-cpu-baseline --no-file-to-file"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_bt -s 0 -c 1 -f -o gpurun_out/r2c_bt $CMD > gpurun_out/ncu_bt.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bt.log
