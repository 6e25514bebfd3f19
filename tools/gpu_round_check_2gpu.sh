#!/bin/bash
# 2-GPU leg of the round check: sharded parity (both exchanges + the 80k rows test), then the 2-GPU bench lines.
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_round_check_2gpu.sh'
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/multi_2
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
lap "pytest shard"
timeout 600 python -m pytest tests/test_shard_gpu.py -m gpu -x -q --durations=5 > $O/pytest_shard.log 2>&1; echo "exit $?" >> $O/pytest_shard.log; tail -4 $O/pytest_shard.log
lap "10k sym"
timeout 300 $TR --master-port 29711 bench.py --gpus 2 --no-cpu-baseline > $O/bench_10k_sym.json 2> $O/bench_10k_sym.err
lap "10k sym EKF_ASYNC_X=0"
EKF_ASYNC_X=0 timeout 300 $TR --master-port 29712 bench.py --gpus 2 --no-cpu-baseline > $O/bench_10k_sym_sync.json 2> $O/bench_10k_sym_sync.err
lap "mc"
timeout 300 $TR --master-port 29713 bench.py --gpus 2 --workload mc --no-cpu-baseline > $O/bench_mc.json 2> $O/bench_mc.err
lap "done"
for f in bench_10k_sym bench_10k_sym_sync bench_mc; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-28s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  exch %s clocks %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        d.get('exchange'), (d.get('clocks') or {}).get('reasons')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e)
PY
done
