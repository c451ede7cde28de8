#!/bin/bash
# Scaling round on one 8-GPU box: C5 (4K, 64 spp progressive, 16.8 M tris, row strips) at N = 1, 2, 4, 8 and the C2 frame at N = 8.
# usage: gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_scale.sh [tag]'
TAG=${1:-sc}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/${TAG}_smi.txt 2>&1
run() { # name, N, args...
  name=$1; N=$2; shift; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" \
      > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.err
  echo "$name exit $?"; python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_${name}.json").read().strip().splitlines()[-1])
    print("$name", "N", d["n_gpus"], "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["config"].get("stages_ms", d["config"].get("ms_per_sample")))
except Exception as e:
    print("$name: no line", e)
PY
}
for n in 8 4 2 1; do run c5_n$n $n --workload C5 --steps 3 --warmup 3 --spp 64; done
run c2_samples_n8 8 --steps 20 --warmup 5
run c2_rows_n8 8 --steps 20 --warmup 5 --mode rows
run c2_samples_n4 4 --steps 20 --warmup 5
timeout 300 python -m pytest tests/test_parallel_gpu.py -x -q -k nccl > $OUT/${TAG}_pytest.log 2>&1; tail -2 $OUT/${TAG}_pytest.log
