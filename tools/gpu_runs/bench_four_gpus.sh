#!/bin/bash
# the N = 2 bench as the driver launches it (reference arm first), then N = 1 default
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 3 --warmup 3 --impl reference > gpurun_out/c60_ref_n4.json 2> gpurun_out/c60_ref_n4.err; echo "ref n2 rc=$?"; tail -c 600 gpurun_out/c60_ref_n4.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/c60_n4.json 2> gpurun_out/c60_n4.err; echo "n2 rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/c60_n4.json").read().strip().split("\n")[-1])
print(d["n_gpus"], round(d["ms_per_step"],2), round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["frac"], d.get("scaling"))
print(json.dumps(d.get("file_to_file")))
PY
tail -3 gpurun_out/c60_n4.err
