#!/bin/bash
# A/B of alternative builds: ab_so.sh v2 v4 ...  (orb-slam2-annotation_b200/liborbb200_<v>.so), then the default build
cd orb-slam2-annotation_b200
cp liborbb200.so /tmp/orig.so
for v in "$@" orig; do
  if [ $v = orig ]; then cp /tmp/orig.so liborbb200.so; else cp liborbb200_$v.so liborbb200.so; fi
  cd ..
  python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -1
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2>gpurun_out/bench_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_$v.json'));print('$v value',round(d['value']),'e2e',round(d['e2e']['value']),'share',d['roofline']['step_share'])"
  cd orb-slam2-annotation_b200
done
