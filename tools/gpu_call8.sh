#!/bin/bash
# Round 2, GPU call 8 (8 GPUs): bench at N = 8 and N = 4 (10k + the 80k block), sharded parity at world 4 / 8, multi-device handle at 4 / 8.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline8.txt; }
nvidia-smi -L > $O/smi8.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
lap "bench --gpus 8"
timeout 600 $TR --nproc-per-node 8 --master-port 29731 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_g8.json 2> $O/bench_g8.err; tail -2 $O/bench_g8.err
lap "bench --gpus 4"
timeout 600 $TR --nproc-per-node 4 --master-port 29732 bench.py --gpus 4 --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_g4.json 2> $O/bench_g4.err; tail -2 $O/bench_g4.err
lap "bench --gpus 8, lazy off / defer 8"
EKF_LAZY_FLUSH=0 timeout 300 $TR --nproc-per-node 8 --master-port 29733 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline --also-80k 0 --parity-steps 0 --sustained 0 > $O/bench_g8_lazy0.json 2> $O/bench_g8_lazy0.err
timeout 300 $TR --nproc-per-node 8 --master-port 29734 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline --also-80k 0 --parity-steps 0 --sustained 0 --defer 8 > $O/bench_g8_defer8.json 2> $O/bench_g8_defer8.err
python - <<'PY'
import json
for f in ['bench_g8','bench_g4','bench_g8_lazy0','bench_g8_defer8']:
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1]); r=d['roofline']
        print(f, 'value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f xch %.4f parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d['exchange']['avg_ms'], d.get('parity')))
        if 'scaling_80k' in d:
            b=d['scaling_80k']; print('   80k', b.get('value'), b.get('ms_per_step'), (b.get('roofline') or {}).get('frac'), b.get('parity'), b.get('error'))
    except Exception as e: print(f, 'ERR', e, open('gpurun_out/%s.err'%f).read()[-500:])
PY
lap "multi-device handle tests (2/4/8)"
EKF_XCHG_TIMEOUT_MS=5000 timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu > $O/pytest_multi_g8.log 2>&1; echo "exit $?" >> $O/pytest_multi_g8.log; tail -5 $O/pytest_multi_g8.log
lap "sharded tests (world 4, 8: p2p; 80k at 8)"
EKF_XCHG_TIMEOUT_MS=5000 timeout 1200 python -m pytest tests/test_shard_gpu.py -q -m gpu -s -k "(p2p and (4 or 8)) or 80k" > $O/pytest_shard_g8.log 2>&1; echo "exit $?" >> $O/pytest_shard_g8.log; grep -E "rank 0|passed|failed|exit|rror" $O/pytest_shard_g8.log | tail -40
lap done
