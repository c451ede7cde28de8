#!/bin/bash
# Multi-GPU round on N GPUs of one box: NCCL parity test + bench lines (sample-partitioned, row bands with BVH rebuild / broadcast, C5).
# usage: gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_multi.sh N [tag]'
N=${1:-2}; TAG=${2:-mg}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $OUT/${TAG}_smi.txt 2>&1
nvidia-smi topo -m >> $OUT/${TAG}_smi.txt 2>&1
timeout 600 python -m pytest tests/test_parallel_gpu.py -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log; tail -3 $OUT/${TAG}_pytest.log
run() { # name, args...
  name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" \
      > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.err
  echo "$name exit $?"; tail -c 400 $OUT/${TAG}_${name}.err; python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_${name}.json").read().strip().splitlines()[-1])
    print("$name", "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["config"].get("stages_ms", d["config"].get("ms_per_sample")))
except Exception as e:
    print("$name: no line", e)
PY
}
run samples --steps 20 --warmup 5
run rows_rebuild --steps 20 --warmup 5 --mode rows --bvh rebuild
run rows_broadcast --steps 20 --warmup 5 --mode rows --bvh broadcast
run c5_rebuild --workload C5 --steps 3 --warmup 3 --spp 16 --bvh rebuild
run c5_broadcast --workload C5 --steps 3 --warmup 3 --spp 16 --bvh broadcast
