#!/bin/bash
mkdir -p gpurun_out
BT2GPU_TRACE=1 python bench.py --no-cpu-baseline --no-file-to-file --steps 3 --warmup 3 > gpurun_out/e2e_trace.json 2> gpurun_out/e2e_trace.err
grep "bt2gpu\]" gpurun_out/e2e_trace.err | tail -12
for v in "BT2GPU_PIPE=524288 BT2GPU_PIPE_RAMP=8" "BT2GPU_PIPE=524288 BT2GPU_PIPE_RAMP=4" "BT2GPU_PIPE=393216 BT2GPU_PIPE_RAMP=6" "BT2GPU_PIPE=524288 BT2GPU_PIPE_RAMP=8 BT2GPU_PIPE_TAIL=1" "BT2GPU_PIPE=480000 BT2GPU_PIPE_RAMP=12"; do
  env $v python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('$v', round(d['ms_per_step'],2), round(d['e2e']['value']))"
done
