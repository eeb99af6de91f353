#!/bin/bash
# confirmation + evidence run: GPU suite, smoke, bench (both arms), secondary rows, launch list of bench.py, ncu full of the level launches
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench rc=$?
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print('value',round(d['value']),'e2e',round(d['e2e']['value']),'share',d['roofline']['step_share'],'frac',round(d['roofline']['frac'],4),'ham',round(d['hamming']['value'],1))"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref rc=$?
timeout 600 python tools/bench_rows.py > gpurun_out/rows.json 2> gpurun_out/rows.err; echo rows rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo launchlist rc=$?
bash tools/gpu_ncu_level512.sh
