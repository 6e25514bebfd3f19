#!/bin/bash
# Round 2, GPU call 18 (1 GPU): k_batch_front (linearisation beside gather + pending correction): bitwise test, suite, bench, e2e.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline18.txt; }
lap "front kernel tests"
timeout 300 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "front or continuous or decided" > $O/pytest_front.log 2>&1; echo "exit $?" >> $O/pytest_front.log; tail -12 $O/pytest_front.log
lap "bench"
B="--no-cpu-baseline --steps 100 --no-streaming-variant --sustained 0"
timeout 300 python bench.py $B > $O/b18_10k.json 2> $O/b18_10k.err
EKF_BATCH_FRONT=0 timeout 300 python bench.py $B --parity-steps 0 > $O/b18_10k_nofront.json 2> $O/b18_10k_nofront.err
for f in b18_10k b18_10k_nofront; do
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
EKF_PROFILE_GAPS=1 EKF_SHADOW=1 timeout 300 python tools/e2e_breakdown.py 10000 200 measure 2>&1 | grep -v "last 8"
EKF_BATCH_FRONT=0 EKF_PROFILE_GAPS=1 EKF_SHADOW=1 timeout 300 python tools/e2e_breakdown.py 10000 200 measure 2>&1 | grep -v "last 8"
lap "full suite"
timeout 1500 python -m pytest tests/ -x -q -m gpu > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -5 $O/pytest_gpu_full.log
lap done
