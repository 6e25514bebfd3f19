#!/bin/bash
# Round 2, GPU call 20 (N GPUs, N = $1): sharded parity at world N (peer-memory exchange), bench lazy vs continuous flush on slabs.
N=${1:-8}
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/multi_$N
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline20.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
lap "pytest shard world $N p2p"
timeout 600 python -m pytest tests/test_shard_gpu.py -m gpu -x -q -k "test_sharded_filter_matches_oracle_and_single_gpu and $N-p2p" > $O/pytest_shard20.log 2>&1; echo "exit $?" >> $O/pytest_shard20.log; tail -4 $O/pytest_shard20.log
B="--no-cpu-baseline --steps 100 --no-streaming-variant --sustained 0"
lap "lazy (default)"; timeout 500 $TR --master-port 29711 bench.py --gpus $N $B > $O/b20_lazy.json 2> $O/b20_lazy.err
lap "continuous flush on slabs"; EKF_SHADOW_SLABS=1 timeout 400 $TR --master-port 29713 bench.py --gpus $N $B --also-80k 0 > $O/b20_shadow.json 2> $O/b20_shadow.err
lap "done"
for f in b20_lazy b20_shadow; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-22s N=%d value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  exch %s parity %s' % (
        sys.argv[1].split('/')[-1], d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        (d.get('exchange') or {}).get('avg_ms'), (d.get('parity') or {}).get('max_rel_P')))
    s = d.get('scaling_80k')
    if s: print('   80k: value %.4g ms/step %.3f frac %.3f parity %s' % (s.get('value', 0), s.get('ms_per_step', 0), s.get('roofline', {}).get('frac', 0), (s.get('parity') or {}).get('max_rel_P') or s.get('error')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e, open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
