#!/bin/bash
# last 1-GPU call of the round: full GPU suite + the default bench line of the final code (denoiser fast paths)
OUT=gpurun_out; mkdir -p $OUT
( time timeout 200 python -m pytest tests -x -q -m gpu > $OUT/r2f_pytest.log 2>&1 ) 2>&1 | grep real; tail -2 $OUT/r2f_pytest.log
( time timeout 150 python bench.py > $OUT/r2f_bench.json 2> $OUT/r2f_bench.err ) 2>&1 | grep real; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2f_bench.json"))
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), {k:round(v,3) for k,v in d["config"]["stages_ms"].items()})
print("roofline_denoise", d.get("roofline_denoise")); print("sha1", d.get("last_frame_sha1")); print("c4", json.dumps(d.get("other_configs",{}).get("c4")))
PY
true
