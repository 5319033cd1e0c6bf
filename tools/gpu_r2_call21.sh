#!/bin/bash
mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/c21_bench.json 2> gpurun_out/c21_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c21_bench.json").read().strip().split("\n")[-1])
f=d.get("file_to_file") or {}
print(round(d["ms_per_step"],2), round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
print(json.dumps(f))
PY
CMD="python bench.py --reads 300000 --steps 1 --warmup 1 --no-cpu-baseline --no-file-to-file"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fill -s 0 -c 1 -f -o gpurun_out/r2b_fill $CMD > gpurun_out/ncu_fill.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_fill.log
