#!/bin/bash
# ncu --set full of one frame's kernels in the REFERENCE layout (parity mode), after a plain run of the same command
OUT=gpurun_out; mkdir -p $OUT
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernels --no-strong --no-configs --layout reference"
timeout 300 python bench.py $ARGS > $OUT/r2ref_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pt_|k_build_blas|k_dn_' -s 57 -c 19 \
    -o $OUT/r2ref_prof -f python bench.py $ARGS > $OUT/r2ref_ncu2.log 2>&1
echo "ncu full exit $?"
