#!/bin/bash
# What the driver's scaling step does: the default bench at N = 1, 2, 4, 8 back to back on one 8-GPU box (+ the NCCL parity tests).
# usage: gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_driver_scale.sh [tag]'
TAG=${1:-ds}; OUT=gpurun_out; mkdir -p $OUT
for N in 1 2 4 8; do
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-ref-kernels > $OUT/${TAG}_n1.json 2> $OUT/${TAG}_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 \
        > $OUT/${TAG}_n$N.json 2> $OUT/${TAG}_n$N.err
  fi
  echo "N=$N exit $?"; python - <<PY
import json
try:
    lines = open("$OUT/${TAG}_n$N.json").read().strip().splitlines()
    d = json.loads(lines[-1])
    print("N", d["n_gpus"], "stdout lines", len(lines), "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["config"]["stages_ms"])
except Exception as e:
    print("no line", e)
PY
done
timeout 300 python -m pytest tests/test_parallel_gpu.py -x -q > $OUT/${TAG}_pytest.log 2>&1; tail -2 $OUT/${TAG}_pytest.log
