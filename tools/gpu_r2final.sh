#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/r2_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/r2_pytest.log
timeout 900 python bench.py > $OUT/r2_bench.json 2> $OUT/r2_bench.err; echo "bench exit $?"
timeout 400 python bench.py --layout reference --no-cpu-baseline --no-ref-kernels --no-strong --no-configs > $OUT/r2_bench_reflayout.json 2> $OUT/r2_bench_reflayout.err; echo "bench ref-layout exit $?"
python - <<'PY'
import json
for f in ("r2_bench","r2_bench_reflayout"):
    d=json.load(open(f"gpurun_out/{f}.json"))
    print(f, "value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()}, {k:round(v,3) for k,v in d["roofline_pathtrace_stage"]["kernels_ms"].items()})
    print("   roofline", d["roofline"]["bound"], d["roofline"]["achieved"], d["roofline"]["peak"], d["roofline"]["frac"], "clocks", d["clocks"])
PY
