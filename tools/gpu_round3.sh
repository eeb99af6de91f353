#!/bin/bash
# GPU suite, smoke, bench (both arms), ncu full of the brute-force tensor-core kernel (one launch of a 2 M-row scan)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench rc=$?
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print('value',round(d['value']),'e2e',round(d['e2e']['value']),'share',d['roofline']['step_share'],'frac',round(d['roofline']['frac'],4),'ham',round(d['hamming']['value'],1),d['hamming']['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref rc=$?
timeout 300 python tools/prof_extract.py 1 --match > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bf_top2_tcn -c 1 -o gpurun_out/prof_bf_tc python tools/prof_extract.py 1 --match > gpurun_out/ncu_tc.log 2>&1
echo "ncu rc=$?"
