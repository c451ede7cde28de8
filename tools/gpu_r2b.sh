#!/bin/bash
# round 2, call B: full GPU suite, bench both layouts, ncu launch list + --set full of the wide-layout kernels
TAG=${1:-r2b}
OUT=gpurun_out; mkdir -p $OUT
if [ "$2" != "skiptests" ]; then
timeout 1500 python -m pytest tests -x -q -m gpu --durations=15 -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
grep -E "passed|failed|Error|differ|wide nodes" $OUT/${TAG}_pytest.log | tail -30
fi
timeout 400 python bench.py --steps 20 --warmup 5 --layout wide > $OUT/${TAG}_bench_wide.json 2> $OUT/${TAG}_bench_wide.err; echo "bench wide exit $?"
timeout 400 python bench.py --steps 20 --warmup 5 --layout reference --no-cpu-baseline --no-ref-kernels > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "bench ref exit $?"
python - <<PY
import json
for f in ("wide","ref"):
    try:
        d=json.load(open(f"gpurun_out/${TAG}_bench_{f}.json"))
        print(f, round(d["value"],1), round(d["ms_per_step"],3), round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()}, {k:round(v,3) for k,v in d["roofline_pathtrace_stage"]["kernels_ms"].items()}, round(d["roofline"]["node_visits_per_ray"],1), round(d["roofline"]["leaf_tests_per_ray"],1))
        print(d.get("microbench"))
    except Exception as e: print(f, "ERR", e)
PY
tail -3 $OUT/${TAG}_bench_wide.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernels --layout wide > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pt_trace|k_pt_primary|k_wide_' -s 12 -c 8 \
    -o $OUT/${TAG}_prof -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernels --layout wide > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
ls -la $OUT | tail -12
