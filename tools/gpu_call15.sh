#!/bin/bash
# Round 2, GPU call 15 (2 GPUs): continuous flush on SYM slabs: single-GPU regression of the strip kernels, sharded parity
# (both exchanges, 80k rows, 10k x 40 steps with the continuous flush), multi-device handle, then the 2-GPU bench lines.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/multi_2
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline15.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
lap "single-GPU continuous-flush tests"
timeout 300 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "continuous or lazy or decided_on_device or x_read_back" > $O/pytest_cf.log 2>&1; echo "exit $?" >> $O/pytest_cf.log; tail -4 $O/pytest_cf.log
lap "pytest shard + multi"
timeout 1500 python -m pytest tests/test_shard_gpu.py tests/test_multi_gpu.py -m gpu -x -q --durations=6 > $O/pytest_shard.log 2>&1; echo "exit $?" >> $O/pytest_shard.log; tail -30 $O/pytest_shard.log
lap "10k sym, 2 GPUs"
timeout 400 $TR --master-port 29711 bench.py --gpus 2 --no-cpu-baseline --steps 100 > $O/bench_10k_sym.json 2> $O/bench_10k_sym.err
lap "10k sym, 2 GPUs, EKF_SHADOW=0"
EKF_SHADOW=0 timeout 400 $TR --master-port 29712 bench.py --gpus 2 --no-cpu-baseline --steps 100 --also-80k 0 > $O/bench_10k_sym_noshadow.json 2> $O/bench_10k_sym_noshadow.err
lap "done"
for f in bench_10k_sym bench_10k_sym_noshadow; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-28s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  exch %s parity %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        (d.get('exchange') or {}).get('avg_ms'), d.get('parity')))
    s = d.get('scaling_80k')
    if s: print('   80k: value %.4g ms/step %.3f frac %.3f parity %s' % (s.get('value', 0), s.get('ms_per_step', 0), s.get('roofline', {}).get('frac', 0), s.get('parity') or s.get('error')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e, open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
