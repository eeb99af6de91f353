#!/bin/bash
# One driver for the GPU-box runs (use as: gpurun -- 'bash tools/gpu.sh <mode> [args]').  Every profiler pass comes AFTER a plain run of the
# same command has exited 0; numbers printed under ncu are never bench values.
#   quick                          parity tests + bench line (no profiler, no CPU baseline)
#   round                          round-end confirmation: GPU suite, smoke(), bench (both arms)
#   ab VAR v1 v2 ...               parity tests once, then the bench line for each value of an environment knob
#   launchlist                     ncu launch list (gpu__time_duration) of bench.py itself
#   ncu-level [frames] [skip] [n]  ncu --set full of level3_kernel launches of a C1 batch (default 512 frames, skip 8, 2 launches)
#   ncu-kernel REGEX SKIP N [ENV]  ncu --set full of one kernel family of a 512-frame C1 batch
#   ncu-ham                        ncu --set full of the brute-force scan kernel (ORBB_BF_VARIANT from the environment) on 2000 x 10 M
#   ncu-match                      ncu --set full of the guided-matching kernels at the C2 shape
#   rows                           tools/bench_rows.py (secondary rows) + its ncu launch list
#   pair [reps]                    matcher tests on the stereo / projection paths + tools/pair_timing.py (per-call medians of one KITTI pair)
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
summary() { python - "$1" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'share', {k: round(x, 4) for k, x in d['roofline']['step_share'].items()},
      'frac', round(d['roofline']['frac'], 4), 'ham', round(d['hamming']['value'], 1), round(d['hamming']['roofline']['frac'], 3))
PY
}
mode=$1; shift
case $mode in
  quick)
    timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
    timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; summary gpurun_out/bench.json;;
  round)
    timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
    timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; summary gpurun_out/bench.json
    timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?";;
  ab)
    timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
    VAR=$1; shift
    for v in "$@"; do
      env $VAR=$v timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "bench $VAR=$v rc=$?"
      summary gpurun_out/bench_$v.json
    done;;
  launchlist)
    timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"
    timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_bench.log 2>&1
    echo "ncu launches rc=$?";;
  ncu-level)
    F=${1:-512}; S=${2:-8}; N=${3:-2}
    timeout 300 python tools/prof_extract.py 2 --frames $F > gpurun_out/plain.log 2>&1 &&
    timeout 1500 $NCU -k regex:level3_kernel -s $S -c $N -o gpurun_out/prof_level python tools/prof_extract.py 2 --frames $F > gpurun_out/ncu_full.log 2>&1
    echo "ncu full rc=$?";;
  ncu-kernel)
    timeout 300 env $4 python tools/prof_extract.py 2 --frames 512 > gpurun_out/plain.log 2>&1 &&
    timeout 1200 env $4 $NCU -k regex:$1 -s $2 -c $3 -o gpurun_out/prof_$1 python tools/prof_extract.py 2 --frames 512 > gpurun_out/ncu_full.log 2>&1
    echo "ncu full rc=$?";;
  ncu-ham)
    V=${ORBB_BF_VARIANT:-4}
    ORBB_BF_VARIANT=$V timeout 120 python tools/ham_quick.py > gpurun_out/ham_plain.log 2>&1 &&
    ORBB_BF_VARIANT=$V timeout 900 $NCU -k regex:bf_top2_tc -s 2 -c 1 -o gpurun_out/prof_ham_v$V python tools/ham_quick.py > gpurun_out/ncu_ham.log 2>&1
    echo "ncu rc=$?"; cat gpurun_out/ham_plain.log;;
  ncu-match)
    timeout 300 python tools/match_workload.py 3 > gpurun_out/match_plain.log 2>&1 &&
    timeout 900 $NCU -k regex:'window_search|resolve_projection|stereo_match|stereo_median|grid_|excl_scan' -s 13 -c 16 -o gpurun_out/prof_match python tools/match_workload.py 3 > gpurun_out/ncu_match.log 2>&1
    echo "ncu rc=$?"; cat gpurun_out/match_plain.log;;
  rows)
    timeout 600 python tools/bench_rows.py > gpurun_out/rows.json 2> gpurun_out/rows.err || { tail -5 gpurun_out/rows.err; exit 1; }
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/rows_launches.csv python tools/bench_rows.py > gpurun_out/rows_ncu.log 2>&1
    echo "ncu rc=$?";;
  pair)
    timeout 600 python -m pytest tests/test_gpu_match.py -x -q -k 'stereo or projection or host_form or tracked' 2>&1 | tail -4
    timeout 300 python tools/pair_timing.py ${1:-200} > gpurun_out/pair_timing.json 2> gpurun_out/pair_timing.err; echo "pair rc=$?"; cat gpurun_out/pair_timing.json; tail -3 gpurun_out/pair_timing.err;;
  *) sed -n 2,15p $0; exit 1;;
esac
