#!/bin/bash
# secondary-row measurements: plain run first, then the ncu launch list of the same command (kernel durations, cold-cache)
set -x
timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q -k "color or colour or cvt" 2>&1 | tail -3
timeout 600 python tools/bench_rows.py > gpurun_out/rows.json 2> gpurun_out/rows.err || { tail -5 gpurun_out/rows.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/rows_launches.csv python tools/bench_rows.py > gpurun_out/rows_ncu.log 2>&1
echo ncu rc=$?
