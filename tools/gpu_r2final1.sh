#!/bin/bash
# final 1-GPU call of the round: full GPU suite, smoke, ncu launch list + --set full of one frame (both after plain runs), bench lines
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -x -q -m gpu > $OUT/r2_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/r2_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
bash tools/gpu_round2.sh r2c > $OUT/r2c_round.log 2>&1; tail -3 $OUT/r2c_round.log
timeout 400 python bench.py --layout reference --no-cpu-baseline --no-ref-kernels --no-strong --no-configs > $OUT/r2_bench_reflayout.json 2> $OUT/r2_bench_reflayout.err; echo "bench ref-layout exit $?"
