#!/bin/bash
# 2-GPU check of the direct (unpacked) halo / frame exchanges: multi-GPU tests + the rows-mode strong-scaling numbers
OUT=gpurun_out; mkdir -p $OUT
( time timeout 90 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > $OUT/r2d_pytest.log 2>&1 ) 2>&1 | grep real; tail -3 $OUT/r2d_pytest.log
( time timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --no-configs --steps 30 --warmup 5 > $OUT/r2d_bench_n2.json 2> $OUT/r2d_bench_n2.err ) 2>&1 | grep real
tail -3 $OUT/r2d_bench_n2.err | cut -c1-300
python - <<PY
import json
d=json.load(open("gpurun_out/r2d_bench_n2.json"))
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1))
print("strong:", json.dumps({k:{kk:vv for kk,vv in v.items() if kk!="workload"} for k,v in d["strong"].items()}))
PY
true
