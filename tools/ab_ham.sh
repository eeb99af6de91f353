#!/bin/bash
# A/B of alternative builds on the Hamming scan only: ab_ham.sh v1 v2 ... (orb-slam2-annotation_b200/liborbb200_<v>.so), then the default
# build; ORBB_BF_VARIANT from the environment (default 4)
cd orb-slam2-annotation_b200
cp liborbb200.so /tmp/orig.so
for v in "$@" orig; do
  case $v in
    orig) cp /tmp/orig.so liborbb200.so;;
    *) cp liborbb200_$v.so liborbb200.so;;
  esac
  echo "== $v"
  (cd .. && ORBB_BF_VARIANT=${ORBB_BF_VARIANT:-4} timeout 120 python tools/ham_quick.py 2>&1 | grep -v "variant 1:")
done
