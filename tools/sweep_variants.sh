#!/bin/bash
# Sweep the downdate kernel variants (EKF_DOWNDATE_VARIANT) and deferral depths; one JSON line per run.
out=${1:-gpurun_out/sweep.jsonl}
: > $out
for d in ${DEFERS:-1 2 4 8}; do
  for v in ${VARIANTS:-0 1 2 3 4 5 6 7}; do
    echo "{\"defer\": $d, \"variant\": $v}" >> $out
    EKF_DOWNDATE_VARIANT=$v timeout 60 python bench.py --defer $d --steps 10 --warmup 3 --no-cpu-baseline ${EXTRA} >> $out 2>&1
  done
done
