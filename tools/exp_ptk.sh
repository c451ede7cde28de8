#!/bin/bash
# per-kernel times of one path-traced frame (ncu serialised durations, us)
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ptk.csv python tools/exp_pt_kernels.py "$@" > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/ptk.csv")) if len(r)>5 and r[0].isdigit()]
ks=[(r[4].split("(")[0], float(r[-1])) for r in rows]
print([(k[:16], round(v/1000,1)) for k,v in ks[-11:]])
PY
