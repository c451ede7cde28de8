for ml in 1024 512 256 128; do
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tail_$ml.csv python tools/exp_pt_kernels.py max_loop=$ml > gpurun_out/tail_$ml.log 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/tail_$ml.csv")) if len(r)>5 and r[0].isdigit()]
ks=[(r[4].split("(")[0], float(r[-1])) for r in rows]
# last frame: take last 12 kernels
print("max_loop=$ml", [(k[:14], round(v/1000,1)) for k,v in ks[-11:]])
PY
done
DIAG=1 python tools/exp_pt_kernels.py 2>&1 | grep round
