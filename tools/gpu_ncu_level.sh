#!/bin/bash
# ncu full capture of the level kernel (levels 0..2 of a 64-frame C1 batch) after a plain run of the same command
mkdir -p gpurun_out
timeout 300 python tools/prof_extract.py 2 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:level3_kernel -s 8 -c 2 -o gpurun_out/prof_level python tools/prof_extract.py 2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
