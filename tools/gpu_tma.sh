#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest(tma) rc=$?"; tail -4 gpurun_out/pytest_gpu.log
ORBB_NO_TMA=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_notma.log 2>&1; echo "pytest(no tma) rc=$?"; tail -4 gpurun_out/pytest_gpu_notma.log
for v in 0 1; do
  if [ $v = 1 ]; then export ORBB_NO_TMA=1; else unset ORBB_NO_TMA; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "bench notma=$v rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$v.json'))
print('notma=$v value',round(d['value']),'e2e',round(d['e2e']['value']),'share',{k:round(x,3) for k,x in d['roofline']['step_share'].items()},'frac',round(d['roofline']['frac'],4))
PY
done
