#!/bin/bash
# tests + stage timing + per-kernel ncu metrics of the path-trace kernels (quick iteration loop)
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python tools/exp_stages.py "$@"
N=2 python tools/exp_stages.py "$@" > gpurun_out/p.log 2>&1 && N=2 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'k_pt_|k_pathtrace' -c 13 --csv --log-file gpurun_out/wf_launches.csv python tools/exp_stages.py "$@" > gpurun_out/p2.log 2>&1
