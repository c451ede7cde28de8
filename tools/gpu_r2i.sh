#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
( time timeout 900 python bench.py > $OUT/r2i_bench.json 2> $OUT/r2i_bench.err ) 2>&1 | grep real; echo "bench exit $?"
tail -3 $OUT/r2i_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2i_bench.json"))
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()}, d["last_frame_sha1"])
print("roofline", json.dumps(d["roofline"])[:1500])
print("strong", json.dumps(d["strong"])[:600])
print("configs", json.dumps(d["other_configs"])[:2500])
print("cpu", d["cpu_baseline"])
PY
( time timeout 600 python bench.py --impl reference --steps 5 --warmup 1 ) 2>&1 | tail -5 | cut -c1-600
