#!/bin/bash
# A/B of alternative builds: ab_so.sh v2 v4 ...  (orb-slam2-annotation_b200/liborbb200_<v>.so), then the default build; a name of the
# form ENV=val runs the default build with that environment variable
cd orb-slam2-annotation_b200
cp liborbb200.so /tmp/orig.so
for v in "$@" orig; do
  E=""
  case $v in
    orig) cp /tmp/orig.so liborbb200.so;;
    *=*) cp /tmp/orig.so liborbb200.so; E=$v;;
    *) cp liborbb200_$v.so liborbb200.so;;
  esac
  cd ..
  env $E python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -1
  env $E python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_ab.json 2>gpurun_out/bench_ab.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_ab.json'));print('$v value',round(d['value']),'e2e',round(d['e2e']['value']),'share',{k:round(x,4) for k,x in d['roofline']['step_share'].items()})"
  cd orb-slam2-annotation_b200
done
