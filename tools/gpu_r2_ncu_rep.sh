#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/scale_check.py --ref-mbp 20 --reads 200000 --sample 1000 --repeats 0.1 --opts '--very-sensitive --dpad 12' --tag rep20 --no-reference --keep"
eval $CMD > gpurun_out/rep20_plain.json 2> gpurun_out/rep20_plain.err &&
eval ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_rep20_launches.csv $CMD > gpurun_out/rep20_ncu.log 2>&1
echo rc=$?; tail -c 600 gpurun_out/rep20_plain.json; python - <<'PY'
import csv, collections
t=collections.Counter(); n=collections.Counter()
rows=list(csv.reader(l for l in open("gpurun_out/r2_rep20_launches.csv") if l.startswith('"')))
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
for r in rows[1:]:
    v=float(r[vi].replace(",","")); u=r[ui]
    v*= {"ns":1e-6,"us":1e-3,"ms":1.0,"s":1e3}.get(u,1e-6)
    k=r[ki].split("(")[0][:40]; t[k]+=v; n[k]+=1
tot=sum(t.values())
for k,v in t.most_common(14): print("%-42s %8.1f ms %5.1f%% %6d launches" % (k,v,100*v/tot,n[k]))
PY
