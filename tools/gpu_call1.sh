#!/bin/bash
# Round 2, GPU call 1: new parity tests first (short), then the K3 sweep of the hybrid kernels, then the full suite.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline.txt; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
lap "hybrid flush bit-identity + boundary tests"
timeout 600 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "hybrid or properties or nonfinite or mutated or failed_measure" > $O/pytest_r2_quick.log 2>&1
echo "exit $?" >> $O/pytest_r2_quick.log; tail -5 $O/pytest_r2_quick.log
lap "sweep rank 32"
: > $O/sweep_r02.jsonl
timeout 300 python tools/sweep_flush.py --defer 16 --out $O/sweep_r02.jsonl - 21 30 31 32 33 34 35 36 37 38 > $O/sweep32.log 2>&1; tail -12 $O/sweep32.log
lap "sweep rank 16"
timeout 300 python tools/sweep_flush.py --defer 8 --out $O/sweep_r02.jsonl - 30 31 32 33 34 35 36 37 38 > $O/sweep16.log 2>&1; tail -11 $O/sweep16.log
lap "sweep rank 8 / rank 2"
timeout 200 python tools/sweep_flush.py --defer 4 --steps 20 --out $O/sweep_r02.jsonl - 30 31 33 36 > $O/sweep8.log 2>&1; tail -5 $O/sweep8.log
timeout 200 python tools/sweep_flush.py --defer 1 --steps 8 --out $O/sweep_r02.jsonl - 30 31 33 36 > $O/sweep2.log 2>&1; tail -5 $O/sweep2.log
lap "sweep FULL storage rank 32 / 16"
timeout 200 python tools/sweep_flush.py --defer 16 --storage full --steps 20 --out $O/sweep_r02.jsonl - 31 33 36 > $O/sweep32f.log 2>&1; tail -4 $O/sweep32f.log
lap "long-horizon tests"
timeout 900 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -s -k "config3 or config2 or mode_b" --durations=10 > $O/pytest_r2_long.log 2>&1
echo "exit $?" >> $O/pytest_r2_long.log; tail -15 $O/pytest_r2_long.log
lap "full -m gpu suite (rest)"
timeout 900 python -m pytest tests/ -q -m gpu --deselect tests/test_round2_gpu.py --durations=10 > $O/pytest_gpu_rest.log 2>&1
echo "exit $?" >> $O/pytest_gpu_rest.log; tail -15 $O/pytest_gpu_rest.log
lap "bench default"
timeout 400 python bench.py > $O/bench_10k.json 2> $O/bench_10k.err; tail -c 1500 $O/bench_10k.json
lap done
