#!/bin/bash
# bench line (with the CPU baseline) + ncu launch list of the profiling workload, after a plain run of the same command
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 300 python tools/prof_extract.py 3 --match > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/prof_extract.py 3 --match > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
