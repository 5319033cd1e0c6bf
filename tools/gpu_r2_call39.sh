#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]))
if f: print("   f2f", round(f["value"]), f["reader_sink_threads"], f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]))
PY
}
for v in pm24 p16d24 p8d16 pm20; do
  BT2GPU_LIB=$PWD/bench_work/libbt2gpu_$v.so python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c39_$v.json 2>/dev/null; show gpurun_out/c39_$v.json
done
for t in "4,12" "6,10" "3,13" "8,16" "5,9"; do
  BT2HOST_STREAM_THREADS=$t python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/c39_t_$t.json 2>/dev/null; show gpurun_out/c39_t_$t.json
done
