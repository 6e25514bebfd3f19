#!/bin/bash
# Round 2, final 1-GPU evidence: smoke, default bench line (both arms), launch list of the default step under ncu.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline_final.txt; }
lap "smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -9 $O/smoke.log
lap "default bench"
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_default.json') if l.startswith('{')][-1]); r=d['roofline']
print('value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f launches %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d.get('gpu_launches'), d.get('parity')))
for kx in ('roofline_in_place_variant', 'roofline_streaming_variant', 'sustained'):
    print(kx, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in d.get(kx, {}).items() if a != 'what'})
print('clocks', d.get('clocks')); print('cpu', {k: v for k, v in d.get('cpu_baseline', {}).items() if k in ('value', 'cores', 'kind', 'extrapolated')}, d.get('cpu_structured', {}).get('value'))
e=json.loads([l for l in open('gpurun_out/bench_ref.json') if l.startswith('{')][-1]); print('reference arm', e.get('value'), e.get('unit'), e.get('config') == d.get('config'))
PY
lap "launch list under ncu"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_r02_final.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_launches.log 2>&1
echo "launch list exit $?"
lap done
