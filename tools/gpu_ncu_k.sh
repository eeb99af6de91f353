#!/bin/bash
# ncu full capture of one kernel family of a 512-frame C1 batch: gpu_ncu_k.sh <regex> <skip> <count> [extra env]
mkdir -p gpurun_out
timeout 300 env $4 python tools/prof_extract.py 2 --frames 512 > gpurun_out/plain.log 2>&1 &&
timeout 1200 env $4 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $3 -o gpurun_out/prof_$1 python tools/prof_extract.py 2 --frames 512 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
