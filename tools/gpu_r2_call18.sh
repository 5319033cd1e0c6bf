#!/bin/bash
# full GPU suite, then the default bench line (with the file-to-file stage breakdown)
mkdir -p gpurun_out
nproc; free -g | head -2
( time python -m pytest tests -m gpu -q -x 2>&1 | tail -8 ) > gpurun_out/c18_pytest.log 2>&1; tail -12 gpurun_out/c18_pytest.log
python bench.py > gpurun_out/c18_bench.json 2> gpurun_out/c18_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c18_bench.json").read().strip().split("\n")[-1])
print(d["ms_per_step"], {k: round(v,2) for k,v in d["phase_ms_per_step"].items()}, d["value"], d["e2e"]["value"], d["roofline"]["frac"])
print(json.dumps(d.get("file_to_file")))
PY
