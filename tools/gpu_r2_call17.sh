#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fill.py tests/test_zz_no_alloc.py tests/test_gpu_parity.py -q 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file-to-file > gpurun_out/c17_bench.json 2> gpurun_out/c17_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c17_bench.json").read().strip().split("\n")[-1])
print(d["ms_per_step"], {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, d["e2e"]["value"], d["roofline"]["frac"])
PY
python tools/scale_check.py --ref-mbp 100 --reads 200000 --sample 1000 --no-reference --tag x > gpurun_out/c17_create.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c17_create.json").read().strip().split("\n")[-1]); print(d["host_index_load_s"], d["create_s"], d["create_info"])
PY
