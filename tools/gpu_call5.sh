#!/bin/bash
# Round 2, GPU call 5 (2 GPUs): lazy flush, multi-device FULL debug, sharded + multi tests, bench N=1 (device 0) and N=2.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline5.txt; }
lap "lazy flush + cross-step + config tests (1 GPU)"
timeout 900 python -m pytest tests/test_round2_gpu.py tests/test_parity_gpu.py -x -q -m gpu -k "lazy or cross_step or config3 or config2 or decoupled or hybrid_flush" > $O/pytest_lazy.log 2>&1; echo "exit $?" >> $O/pytest_lazy.log; tail -6 $O/pytest_lazy.log
lap "debug multi full"
EKF_XCHG_TIMEOUT_MS=5000 timeout 300 python tools/debug_multi_full.py > $O/debug_multi_full.log 2>&1; tail -5 $O/debug_multi_full.log
lap "multi-device handle tests"
EKF_XCHG_TIMEOUT_MS=5000 timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu > $O/pytest_multi_g2.log 2>&1; echo "exit $?" >> $O/pytest_multi_g2.log; tail -5 $O/pytest_multi_g2.log
lap "sharded tests (world 2)"
EKF_XCHG_TIMEOUT_MS=5000 timeout 900 python -m pytest tests/test_shard_gpu.py -x -q -m gpu -s > $O/pytest_shard_g2.log 2>&1; echo "exit $?" >> $O/pytest_shard_g2.log; grep -E "rank 0|passed|failed|exit|rror" $O/pytest_shard_g2.log | tail -30
lap "bench N=1 lazy on / off"
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 > $O/bench_lazy1.json 2> $O/bench_lazy1.err; tail -c 300 $O/bench_lazy1.json
EKF_LAZY_FLUSH=0 timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 --parity-steps 0 > $O/bench_lazy0.json 2> $O/bench_lazy0.err
lap "bench --gpus 2"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_g2.json 2> $O/bench_g2.err
tail -3 $O/bench_g2.err
EKF_LAZY_FLUSH=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29722 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --also-80k 0 --parity-steps 0 > $O/bench_g2_lazy0.json 2> $O/bench_g2_lazy0.err
python - <<'PY'
import json
for f in ['bench_lazy1','bench_lazy0','bench_g2','bench_g2_lazy0']:
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1])
        r=d['roofline']; print(f, 'value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d.get('parity')))
        if 'scaling_80k' in d: print('   80k', d['scaling_80k'].get('value'), d['scaling_80k'].get('parity'))
    except Exception as e: print(f, 'ERR', e)
PY
lap done
