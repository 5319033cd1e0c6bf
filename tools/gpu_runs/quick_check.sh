#!/bin/bash
# parity tests of the kernels + a short bench line
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fill.py tests/test_gpu_parity.py tests/test_long_reads.py tests/test_gpu_fullsize.py tests/test_gpu_vs_reference_binary.py tests/test_kat.py tests/test_fuzz_options.py -q -x -m gpu 2>&1 | tail -2
python bench.py --no-cpu-baseline --no-file-to-file --steps 5 --warmup 3 > gpurun_out/quick_bench.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open("gpurun_out/quick_bench.json").read().strip().split("\n")[-1])
print(round(d["ms_per_step"],2), {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"],3))
PY
