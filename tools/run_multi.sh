#!/bin/bash
# Multi-GPU checks on one box: sharded-filter parity, then the strong-scaling bench lines.
# usage: RUNS="shard 10k_sym 10k_full mc 80k_sym" run_multi.sh N
N=${1:-2}
RUNS=${RUNS:-"shard 10k_sym 10k_full mc 80k_sym"}
out=gpurun_out/multi_${N}
mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for r in $RUNS; do
  case $r in
    shard) timeout 900 python -m pytest tests/test_shard_gpu.py -m gpu -x -q > $out/pytest_shard.log 2>&1; tail -3 $out/pytest_shard.log ;;
    10k_sym) timeout 300 $TR --master-port 29711 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_10k_sym.json 2> $out/bench_10k_sym.err ;;
    10k_full) timeout 300 $TR --master-port 29712 bench.py --gpus $N --steps 20 --warmup 3 --storage full --no-cpu-baseline > $out/bench_10k_full.json 2> $out/bench_10k_full.err ;;
    mc) timeout 300 $TR --master-port 29713 bench.py --gpus $N --workload mc --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_mc.json 2> $out/bench_mc.err ;;
    80k_sym) timeout 420 $TR --master-port 29714 bench.py --gpus $N --landmarks 80000 --steps 5 --warmup 3 --no-cpu-baseline > $out/bench_80k_sym.json 2> $out/bench_80k_sym.err ;;
    80k_full) timeout 420 $TR --master-port 29715 bench.py --gpus $N --landmarks 80000 --storage full --steps 5 --warmup 3 --no-cpu-baseline > $out/bench_80k_full.json 2> $out/bench_80k_full.err ;;
  esac
done
for f in $out/bench_*.json; do echo "== $f"; grep -v NCCL $f | head -c 1500; echo; grep -E "Error|error" ${f%.json}.err | tail -3; done
