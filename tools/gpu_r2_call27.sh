#!/bin/bash
mkdir -p gpurun_out
python tools/f2f_probe.py 1000000 4 > gpurun_out/c27_probe.log 2> gpurun_out/c27_probe.err; cat gpurun_out/c27_probe.log
for t in "5,9" "6,8" "8,8" "4,6" "16,16"; do
BT2HOST_STREAM_THREADS=$t python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/c27_t_$t.json 2>/dev/null
python - "$t" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/c27_t_%s.json"%sys.argv[1]).read().strip().split("\n")[-1]); f=d["file_to_file"]
print(sys.argv[1], "f2f", round(f["value"]), f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
done
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 4000 --long 250,500,750,1000 --opts "--sensitive" --tag c5_long_100mbp > gpurun_out/c27_c5.json 2> gpurun_out/c27_c5.err; echo "c5 rc=$?"; cat gpurun_out/c27_c5.json; tail -3 gpurun_out/c27_c5.err
