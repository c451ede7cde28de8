#!/bin/bash
# final state: ncu launch list of one bench command + ncu --set full of the denoiser's launches of one frame (after a plain run exited 0)
OUT=gpurun_out; mkdir -p $OUT
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernels --no-strong --no-configs"
timeout 60 python bench.py $ARGS > $OUT/r2f_plain.log 2>&1 &&
timeout 70 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2f_launches.csv python bench.py $ARGS > $OUT/r2f_ncu1.log 2>&1
echo "ncu launches exit $?"
timeout 70 ncu --set full --clock-control none --import-source on -k regex:'k_dn_' -s 16 -c 8 -o $OUT/r2f_dn_prof -f python bench.py $ARGS > $OUT/r2f_ncu2.log 2>&1
echo "ncu full exit $?"
ls -la $OUT | tail -4
