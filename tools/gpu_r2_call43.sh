#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
if f: print("   f2f", round(f["value"]), f["reader_sink_threads"], f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
}
python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py -q -x 2>&1 | tail -2
for t in "5,9" "4,6" "4,8" "6,6" "5,5"; do
  BT2HOST_STREAM_THREADS=$t python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/c43_t_$t.json 2>/dev/null; show gpurun_out/c43_t_$t.json
done
