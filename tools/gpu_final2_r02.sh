#!/bin/bash
# Round 2, closing run on HEAD (1 GPU): full -m gpu suite with -x, smoke, default bench line.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline_final2.txt; }
lap "full -m gpu suite, -x"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=4 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -10 $O/pytest_gpu_full.log
lap "smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
lap "default bench"
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_default.json') if l.startswith('{')][-1]); r=d['roofline']
print('value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f launches %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d.get('gpu_launches'), d.get('parity')))
print('in place', d.get('roofline_in_place_variant', {}).get('frac'), 'sustained', d.get('sustained', {}).get('value'), 'clocks', d.get('clocks'))
PY
lap "mc bench"
timeout 300 python bench.py --workload mc --no-cpu-baseline > $O/bench_mc.json 2> $O/bench_mc.err; python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_mc.json') if l.startswith('{')][-1]); r=d['roofline']
print('mc value %.4g e2e %.4g ms/step %.4f frac %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['frac']))
PY
lap done
