#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1])
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
PY
}
for s in 262144 524288 1048576; do
  BT2GPU_SLOTS=$s python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c34_s$s.json 2>/dev/null; show gpurun_out/c34_s$s.json
done
BT2GPU_SLOTS=1048576 BT2GPU_PIPE=524288 python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c34_s1m_p512k.json 2>/dev/null; show gpurun_out/c34_s1m_p512k.json
