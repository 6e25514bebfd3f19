#!/bin/bash
# Round 2, GPU call 3 (2 GPUs): sharded parity incl. cross-step deferral on slabs, single-process multi-device handle, bench at N = 2.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline3.txt; }
nvidia-smi -L > $O/smi3.txt
lap "multi-device handle tests"
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > $O/pytest_multi_g2.log 2>&1; echo "exit $?" >> $O/pytest_multi_g2.log; tail -6 $O/pytest_multi_g2.log
lap "sharded tests (world 2)"
timeout 900 python -m pytest tests/test_shard_gpu.py -x -q -m gpu -s > $O/pytest_shard_g2.log 2>&1; echo "exit $?" >> $O/pytest_shard_g2.log; grep -E "rank 0|passed|failed|exit|Error" $O/pytest_shard_g2.log | tail -25
lap "bench --gpus 2 (default: 10k + 80k block)"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_g2.json 2> $O/bench_g2.err
grep -v NCCL $O/bench_g2.json | tail -c 3000; tail -5 $O/bench_g2.err
lap "bench --gpus 2 defer 8"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29722 bench.py --gpus 2 --steps 20 --warmup 3 --defer 8 --also-80k 0 --no-cpu-baseline --sustained 0 > $O/bench_g2_defer8.json 2> $O/bench_g2_defer8.err
grep -v NCCL $O/bench_g2_defer8.json | tail -c 600
lap "reference arm under torchrun"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29723 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_ref_g2.json 2> $O/bench_ref_g2.err
tail -c 700 $O/bench_ref_g2.json
lap done
