#!/bin/bash
# Round 2, GPU call 19 (2 GPUs): sharded parity with the front kernel on slabs (pending terms) + 2-GPU bench lines.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/multi_2
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline19.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
lap "pytest shard + multi"
timeout 1500 python -m pytest tests/test_shard_gpu.py tests/test_multi_gpu.py -m gpu -x -q --durations=4 > $O/pytest_shard19.log 2>&1; echo "exit $?" >> $O/pytest_shard19.log; tail -12 $O/pytest_shard19.log
B="--no-cpu-baseline --steps 100 --no-streaming-variant --sustained 0"
lap "2 GPUs"; timeout 400 $TR --master-port 29711 bench.py --gpus 2 $B > $O/b19_g2.json 2> $O/b19_g2.err
lap "2 GPUs EKF_BATCH_FRONT=0"; EKF_BATCH_FRONT=0 timeout 400 $TR --master-port 29713 bench.py --gpus 2 $B --also-80k 0 > $O/b19_g2_nofront.json 2> $O/b19_g2_nofront.err
lap "done"
for f in b19_g2 b19_g2_nofront; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-22s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  exch %s parity %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        (d.get('exchange') or {}).get('avg_ms'), (d.get('parity') or {}).get('max_rel_P')))
    s = d.get('scaling_80k')
    if s: print('   80k: value %.4g ms/step %.3f frac %.3f parity %s' % (s.get('value', 0), s.get('ms_per_step', 0), s.get('roofline', {}).get('frac', 0), s.get('parity') or s.get('error')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e, open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
