#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py -x -q -m gpu 2>&1 | tail -2
python tools/exp_wide_ab.py wide_bvh=1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_wide_' -c 9 python tools/exp_wide_ab.py wide_bvh=1 2>&1 | grep -E "gpu__time" | head -9
