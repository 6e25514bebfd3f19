#!/bin/bash
# Round 2, GPU call 17 (1 GPU): k_batch_lin with staged pending terms: full suite, bench 10k / 1k, e2e breakdown.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline17.txt; }
lap "full -m gpu suite, -x"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=3 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -8 $O/pytest_gpu_full.log
lap "bench"
B="--no-cpu-baseline --steps 100 --no-streaming-variant --sustained 0"
timeout 300 python bench.py $B > $O/b17_10k.json 2> $O/b17_10k.err
EKF_SHADOW=0 timeout 300 python bench.py $B --parity-steps 0 > $O/b17_10k_noshadow.json 2> $O/b17_10k_noshadow.err
timeout 300 python bench.py $B --landmarks 1000 --steps 200 --parity-steps 0 > $O/b17_1k.json 2> $O/b17_1k.err
timeout 300 python bench.py $B --defer 8 --parity-steps 0 > $O/b17_10k_d8.json 2> $O/b17_10k_d8.err
for f in b17_10k b17_10k_noshadow b17_1k b17_10k_d8; do
    python - "$O/$f.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
    r = d['roofline']
    print('%-22s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f parity %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], (d.get('parity') or {}).get('max_rel_P')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e, open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
lap "e2e breakdown"
EKF_PROFILE_GAPS=1 timeout 300 python tools/e2e_breakdown.py 10000 200 2>&1 | grep -v "last 8"
lap done
