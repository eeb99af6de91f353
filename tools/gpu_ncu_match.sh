#!/bin/bash
# ncu full capture of the guided-matching kernels at the C2 shape, after a plain run of the same command
mkdir -p gpurun_out
timeout 300 python tools/match_workload.py 3 > gpurun_out/match_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'window_search|resolve_projection|stereo_match|stereo_median|grid_|excl_scan' -s 13 -c 16 -o gpurun_out/prof_match python tools/match_workload.py 3 > gpurun_out/ncu_match.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/match_plain.log; tail -3 gpurun_out/ncu_match.log
