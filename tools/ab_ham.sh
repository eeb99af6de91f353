#!/bin/bash
# A/B of alternative builds on the brute-force top-2 workload: ab_ham.sh v1 v2 ... (orb-slam2-annotation_b200/liborbb200_<v>.so), then the default
cd orb-slam2-annotation_b200
cp liborbb200.so /tmp/orig.so
for v in "$@" orig; do
  if [ $v = orig ]; then cp /tmp/orig.so liborbb200.so; else cp liborbb200_$v.so liborbb200.so; fi
  cd ..
  echo "== $v"; timeout 200 python tools/ham_quick.py 2>&1 | tail -3
  cd orb-slam2-annotation_b200
done
