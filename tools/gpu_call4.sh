#!/bin/bash
# Round 2, GPU call 4 (2 GPUs): multi-device front (lockstep), mode B on slabs, sharded parity again.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline4.txt; }
lap "multi-device handle tests"
EKF_XCHG_TIMEOUT_MS=5000 timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu > $O/pytest_multi_g2.log 2>&1; echo "exit $?" >> $O/pytest_multi_g2.log; tail -12 $O/pytest_multi_g2.log
lap "sharded tests (world 2)"
EKF_XCHG_TIMEOUT_MS=5000 timeout 900 python -m pytest tests/test_shard_gpu.py -x -q -m gpu -s > $O/pytest_shard_g2.log 2>&1; echo "exit $?" >> $O/pytest_shard_g2.log; grep -E "rank 0|passed|failed|exit|Error|rror" $O/pytest_shard_g2.log | tail -30
lap done
