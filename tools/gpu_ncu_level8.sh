#!/bin/bash
# ncu full capture of the 8 level-kernel launches of one 64-frame C1 batch, after a plain run of the same command
mkdir -p gpurun_out
timeout 300 python tools/prof_extract.py 2 > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:level3_kernel -s 8 -c 8 -o gpurun_out/prof_level8 python tools/prof_extract.py 2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
