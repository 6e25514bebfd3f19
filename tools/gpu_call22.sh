#!/bin/bash
# Round 2, GPU call 22 (1 GPU): moving window base (no freeze copy, no per-pass rotation): suite, bench, e2e.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline22.txt; }
lap "continuous-flush tests"
timeout 300 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "front or continuous or decided or x_read" > $O/pytest_cf22.log 2>&1; echo "exit $?" >> $O/pytest_cf22.log; tail -12 $O/pytest_cf22.log
lap "bench"
B="--no-cpu-baseline --steps 400 --no-streaming-variant --sustained 0"
timeout 300 python bench.py $B > $O/b22_10k.json 2> $O/b22_10k.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/b22_10k.json') if l.startswith('{')][-1]); r=d['roofline']
print('value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f launches %s parity %s clocks %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d.get('gpu_launches'), (d.get('parity') or {}).get('max_rel_P'), (d.get('clocks') or {}).get('reasons')))
PY
lap "e2e breakdown"
EKF_PROFILE_GAPS=1 EKF_SHADOW=1 timeout 300 python tools/e2e_breakdown.py 10000 400 measure 2>&1 | grep -v "last 8"
lap "full suite"
timeout 1500 python -m pytest tests/ -x -q -m gpu > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -5 $O/pytest_gpu_full.log
lap done
