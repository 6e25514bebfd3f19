#!/bin/bash
# Round 2, GPU call 6 (1 GPU): full -m gpu suite after the dot3 / lazy / mode-B changes; ncu --set full of k_batch_lin / k_batch_gain.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline6.txt; }
lap "full -m gpu suite"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=8 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -14 $O/pytest_gpu_full.log
lap "ncu lin/gain"
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/plain_lg.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_batch_ -s 4 -c 6 -f -o $O/ncu_batch_lin_gain_r02 \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_lg.log 2>&1
echo "ncu exit $?"
lap done
