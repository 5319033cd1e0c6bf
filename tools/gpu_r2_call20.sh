#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -x 2>&1 | tail -3
python bench.py > gpurun_out/c20_bench.json 2> gpurun_out/c20_bench.err; echo "bench rc=$?"
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1])
f=d.get("file_to_file") or {}
print(sys.argv[1], round(d["ms_per_step"],2), round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3), "| f2f", round(f.get("value",0)), f.get("reader_sink_threads"), f.get("stage_busy_s"))
PY
}
show gpurun_out/c20_bench.json
for t in "4,10" "6,8" "8,8" "16,16" "3,13"; do
  BT2HOST_STREAM_THREADS=$t python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/c20_t_$t.json 2>/dev/null; show gpurun_out/c20_t_$t.json
done
BT2HOST_SAM_WRITE=pwrite python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/c20_pwrite.json 2>/dev/null; show gpurun_out/c20_pwrite.json
