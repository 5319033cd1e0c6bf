#!/bin/bash
mkdir -p gpurun_out
run() { # tag, env...
  tag=$1; shift
  env "$@" python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/c28_$tag.json 2>/dev/null
  python - "$tag" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/c28_%s.json"%sys.argv[1]).read().strip().split("\n")[-1]); f=d["file_to_file"]
print(sys.argv[1], "f2f", round(f["value"]), f["ranges"], f["stage_busy_s"], "devnull", round(f["sam_to_dev_null"]["value"]), f["sam_to_dev_null"]["stage_busy_s"])
PY
}
run base X=1
run nopop BT2HOST_SAM_POPULATE=0
run c96 BT2HOST_STREAM_CHUNK_MB=96
run c128 BT2HOST_STREAM_CHUNK_MB=128
run c48 BT2HOST_STREAM_CHUNK_MB=48
run pipe64k BT2GPU_PIPE=65536
run pipe128k BT2GPU_PIPE=131072
run c32pipe64k BT2HOST_STREAM_CHUNK_MB=32 BT2GPU_PIPE=65536
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 4000 --long 250,500,750,1000 --opts=--sensitive --tag c5_long_100mbp > gpurun_out/c28_c5.json 2> gpurun_out/c28_c5.err; echo "c5 rc=$?"; cat gpurun_out/c28_c5.json; tail -3 gpurun_out/c28_c5.err
