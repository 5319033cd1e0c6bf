#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1])
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]))
PY
}
for v in pm4 pm16 p16d4 t64; do
  BT2GPU_LIB=$PWD/bench_work/libbt2gpu_$v.so python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c38_$v.json 2>/dev/null; show gpurun_out/c38_$v.json
done
python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c38_base.json 2>/dev/null; show gpurun_out/c38_base.json
