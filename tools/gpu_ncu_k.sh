#!/bin/bash
# ncu full capture of one kernel (regex $1, skip $2, count $3) of the 64-frame C1 workload, after a plain run of the same command
mkdir -p gpurun_out
timeout 300 python tools/prof_extract.py 2 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $3 -o gpurun_out/prof_$1 python tools/prof_extract.py 2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
