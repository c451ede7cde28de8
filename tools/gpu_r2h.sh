#!/bin/bash
# 2-GPU call: NCCL tests of the C-level multi-GPU path, bench --gpus 2 (samples default with strong sub-records; rows mode)
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 600 python -m pytest tests/test_multi_gpu.py tests/test_parallel_gpu.py -x -q -m gpu 2>&1 | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $OUT/r2h_bench_n2.json 2> $OUT/r2h_bench_n2.err; echo "bench n2 exit $?"
tail -5 $OUT/r2h_bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --mode rows --no-strong > $OUT/r2h_bench_n2_rows.json 2> $OUT/r2h_bench_n2_rows.err; echo "bench n2 rows exit $?"
tail -5 $OUT/r2h_bench_n2_rows.err
python - <<'PY'
import json
for f in ("r2h_bench_n2","r2h_bench_n2_rows"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, "value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()})
        print("  strong:", json.dumps(d.get("strong"))[:900])
    except Exception as e: print(f, "ERR", e)
PY
