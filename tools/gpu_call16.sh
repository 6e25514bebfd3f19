#!/bin/bash
# Round 2, GPU call 16 (2 GPUs): one-row-per-thread gather / scatter; 1-GPU and 2-GPU bench, default and 2-CTA/SM pass shapes.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/multi_2
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline16.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
lap "shard parity (p2p) + continuous 10k"
timeout 900 python -m pytest tests/test_shard_gpu.py -m gpu -x -q -k "p2p or continuous" > $O/pytest_shard16.log 2>&1; echo "exit $?" >> $O/pytest_shard16.log; tail -4 $O/pytest_shard16.log
B="--no-cpu-baseline --steps 100 --also-80k 0 --no-streaming-variant --sustained 0"
lap "1 GPU"; timeout 300 python bench.py $B > $O/b1.json 2> $O/b1.err
lap "1 GPU v33"; EKF_SHADOW_VARIANT=33 timeout 300 python bench.py $B > $O/b1_v33.json 2> $O/b1_v33.err
lap "2 GPUs"; timeout 400 $TR --master-port 29711 bench.py --gpus 2 $B > $O/b2.json 2> $O/b2.err
lap "2 GPUs v33"; EKF_SHADOW_VARIANT=33 timeout 400 $TR --master-port 29712 bench.py --gpus 2 $B > $O/b2_v33.json 2> $O/b2_v33.err
lap "2 GPUs EKF_SHADOW=0"; EKF_SHADOW=0 timeout 400 $TR --master-port 29713 bench.py --gpus 2 $B > $O/b2_noshadow.json 2> $O/b2_noshadow.err
lap "done"
for f in b1 b1_v33 b2 b2_v33 b2_noshadow; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-22s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  exch %s parity %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        (d.get('exchange') or {}).get('avg_ms'), (d.get('parity') or {}).get('max_rel_P')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e, open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
