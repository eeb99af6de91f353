#!/bin/bash
# round-end confirmation: full GPU suite, smoke, bench (both arms), secondary rows
set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench rc=$?
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref rc=$?
timeout 500 python tools/bench_rows.py > gpurun_out/rows.json 2> gpurun_out/rows.err; echo rows rc=$?
