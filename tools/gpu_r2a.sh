#!/bin/bash
# round 2, call A: full GPU suite with durations, bench line in both layouts
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu --durations=30 -s > $OUT/r2a_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/r2a_pytest.log
tail -45 $OUT/r2a_pytest.log
timeout 400 python bench.py --steps 20 --warmup 5 --layout wide > $OUT/r2a_bench_wide.json 2> $OUT/r2a_bench_wide.err; echo "bench wide exit $?"
timeout 400 python bench.py --steps 20 --warmup 5 --layout reference --no-cpu-baseline --no-ref-kernels > $OUT/r2a_bench_ref.json 2> $OUT/r2a_bench_ref.err; echo "bench ref exit $?"
python - <<'PY'
import json
for f in ("wide","ref"):
    try:
        d=json.load(open(f"gpurun_out/r2a_bench_{f}.json"))
        print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"]["stages_ms"], d["roofline_pathtrace_stage"]["kernels_ms"], d["roofline"]["node_visits_per_ray"], d["roofline"]["leaf_tests_per_ray"])
    except Exception as e: print(f, "ERR", e)
PY
tail -3 $OUT/r2a_bench_wide.err
