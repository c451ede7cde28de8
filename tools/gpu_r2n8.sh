#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
N=${1:-8}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
( time timeout 870 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N > $OUT/r2_bench_n$N.json 2> $OUT/r2_bench_n$N.err ) 2>&1 | grep real; echo "bench n$N exit $?"
tail -4 $OUT/r2_bench_n$N.err | cut -c1-300
true
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_n$N.json"))
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()})
print("strong:", json.dumps({k:{kk:vv for kk,vv in v.items() if kk!="workload"} for k,v in d["strong"].items()}))

PY
