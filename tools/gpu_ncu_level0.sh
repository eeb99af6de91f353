#!/bin/bash
# ncu full capture (with source counters) of the level-0 and level-1 launches of one 512-frame C1 batch, after a plain run
mkdir -p gpurun_out
timeout 300 python tools/prof_extract.py 2 --frames 512 > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:level3_kernel -s 8 -c 2 -o gpurun_out/prof_level0 python tools/prof_extract.py 2 --frames 512 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
