#!/bin/bash
# parity tests once, then the bench line for each value of an env knob:  gpu_ab.sh VAR v1 v2 ...
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "bench $VAR=$v rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$v.json'))
print('$VAR=$v value',round(d['value']),'e2e',round(d['e2e']['value']),'share',{k:round(x,3) for k,x in d['roofline']['step_share'].items()},'frac',round(d['roofline']['frac'],4))
PY
done
