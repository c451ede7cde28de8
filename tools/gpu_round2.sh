#!/bin/bash
# round-2 profile call: ncu launch list + ncu --set full of one whole frame's kernels (after a plain run of the same command exited 0)
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernels --no-strong --no-configs"
timeout 300 python bench.py $ARGS > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv python bench.py $ARGS > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
timeout 300 python bench.py $ARGS > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pt_|k_build_blas|k_wide_|k_dn_' -s 66 -c 22 \
    -o $OUT/${TAG}_prof -f python bench.py $ARGS > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
ls -la $OUT | tail -6
