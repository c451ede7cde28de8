#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py -x -q -m gpu -s 2>&1 | grep -E "passed|failed|differ|Error|assert" | head -20
python tools/exp_wide_ab.py wide_bvh=0
python tools/exp_wide_ab.py wide_bvh=1
python tools/exp_wide_ab.py wide_bvh=2
python tools/exp_wide_ab.py wide_bvh=1 > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_pt_trace|k_pt_primary' -s 4 -c 3 -o gpurun_out/r2e_prof -f python tools/exp_wide_ab.py wide_bvh=1 > gpurun_out/r2e_ncu.log 2>&1
echo ncu exit $?
