#!/bin/bash
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().split("\n")[-1]); f=d.get("file_to_file") or {}
print(sys.argv[1], round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]))
if f: print("   f2f", round(f["value"]), f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]))
PY
}
python -m pytest tests/test_long_reads.py tests/test_gpu_fill.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -x -m gpu 2>&1 | tail -2
for s in 0 262144 131072; do
  BT2GPU_FILL_SPLIT=$s python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/c42_split$s.json 2>/dev/null; show gpurun_out/c42_split$s.json
done
python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/c42_prealloc.json 2>/dev/null; show gpurun_out/c42_prealloc.json
BT2HOST_STREAM_PREALLOC=0 python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/c42_noprealloc.json 2>/dev/null; show gpurun_out/c42_noprealloc.json
python bench.py --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/c42_prealloc2.json 2>/dev/null; show gpurun_out/c42_prealloc2.json
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 4000 --long 250,500,750,1000 --opts=--sensitive --tag c5_long_100mbp > gpurun_out/c42_c5.json 2> gpurun_out/c42_c5.err; echo "c5 rc=$?"; cat gpurun_out/c42_c5.json
