#!/bin/bash
# A/B of K_fill build variants: build/libbt2gpu_<v>.so copied over the package's library in turn
mkdir -p gpurun_out
for v in "$@"; do
  cp build/libbt2gpu_$v.so omp-bowtie2-prime_b200/libbt2gpu.so
  timeout 300 python -m pytest tests/test_gpu_fill.py -x -q > gpurun_out/var_${v}_test.log 2>&1; echo "$v fill test rc=$?"
  timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/var_$v.json").read().strip().split("\n")[-1])
print("$v", "ms/step %.2f" % d["ms_per_step"], {k: round(x, 2) for k, x in d["phase_ms_per_step"].items()}, "e2e %.2f M/s" % (d["e2e"]["value"] / 1e6))
PY
done
