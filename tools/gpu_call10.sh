#!/bin/bash
# Round 2, GPU call 10 (1 GPU): the no-allocation association entry point alone in a fresh process, then the whole
# -m gpu suite with -x (the static-scratch tests are collected last), then the default bench line.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline10.txt; }
lap "static-scratch test alone"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_zz_alone.log 2>&1; echo "exit $?" >> $O/pytest_zz_alone.log; tail -15 $O/pytest_zz_alone.log
lap "full -m gpu suite, -x"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=8 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -14 $O/pytest_gpu_full.log
lap "default bench"
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"; tail -c 1500 $O/bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"; tail -c 600 $O/bench_ref.json
lap done
