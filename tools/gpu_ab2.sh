#!/bin/bash
# bench line for several "VAR=val,VAR2=val2" env combinations (no tests):  gpu_ab2.sh combo1 combo2 ...
mkdir -p gpurun_out
i=0
for combo in "$@"; do
  i=$((i+1))
  env $(echo $combo | tr ',' ' ') timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c$i.json 2> gpurun_out/bench_c$i.err; echo "bench $combo rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_c$i.json'))
print('$combo value',round(d['value']),'e2e',round(d['e2e']['value']),'share',{k:round(x,3) for k,x in d['roofline']['step_share'].items()})
PY
done
