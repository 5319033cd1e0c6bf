#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
if f: print("   f2f", round(f["value"]), f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
}
for v in u2 u4; do
  BT2GPU_LIB=$PWD/bench_work/libbt2gpu_$v.so python -m pytest tests/test_gpu_fill.py tests/test_gpu_parity.py -q -x 2>&1 | tail -1
  BT2GPU_LIB=$PWD/bench_work/libbt2gpu_$v.so python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c32_$v.json 2>/dev/null; show gpurun_out/c32_$v.json
done
python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/c32_pinned.json 2>/dev/null; show gpurun_out/c32_pinned.json
python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/c32_pinned2.json 2>/dev/null; show gpurun_out/c32_pinned2.json
BT2_BENCH_PINNED=0 python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/c32_pageable.json 2>/dev/null; show gpurun_out/c32_pageable.json
