#!/bin/bash
# ncu full capture of the brute-force scan kernel (ORBB_BF_VARIANT from the environment, default 4) on 2000 x 10 M, after a plain run
mkdir -p gpurun_out
V=${ORBB_BF_VARIANT:-4}
ORBB_BF_VARIANT=$V timeout 120 python tools/ham_quick.py > gpurun_out/ham_plain.log 2>&1 &&
ORBB_BF_VARIANT=$V timeout 900 ncu --set full --clock-control none --import-source on -k regex:bf_top2_tc -s 2 -c 1 -o gpurun_out/prof_ham_v$V python tools/ham_quick.py > gpurun_out/ncu_ham.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/ham_plain.log
