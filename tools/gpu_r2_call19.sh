#!/bin/bash
# host stage scaling on the box's cores; the two-stream chunk schedule on/off with several k_bt residencies
mkdir -p gpurun_out
python tests/host_rate.py 1000000 /dev/shm/hr --gpu 2>&1 | tail -12
rm -rf /dev/shm/hr
python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference_binary.py -q -x 2>&1 | tail -3
for v in "0 3" "1 2" "1 3" "1 4" "1 6"; do set -- $v
  BT2GPU_OVERLAP=$1 BT2GPU_BT_BLOCKS=$2 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file-to-file > gpurun_out/c19_b_$1_$2.json 2>/dev/null
  python - "$1" "$2" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/c19_b_%s_%s.json"%(sys.argv[1],sys.argv[2])).read().strip().split("\n")[-1])
print("overlap",sys.argv[1],"btblocks",sys.argv[2],round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["e2e"]["value"]))
PY
done
