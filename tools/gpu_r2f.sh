#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py tests/test_traverse_gpu.py tests/test_pathtrace_gpu.py -x -q -m gpu 2>&1 | tail -3
python tools/exp_wide_ab.py wide_bvh=0
python tools/exp_wide_ab.py wide_bvh=1
python tools/exp_wide_ab.py wide_bvh=1 > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_pt_trace|k_pt_primary' -s 4 -c 3 -o gpurun_out/r2f_prof -f python tools/exp_wide_ab.py wide_bvh=1 > gpurun_out/r2f_ncu.log 2>&1
echo ncu exit $?
