#!/bin/bash
# Round 2, GPU call 14 (1 GPU): continuous flush as shipped: full suite, default bench line, launch list + ncu --set full capture
# of the out-of-place rank-32 pass (each only after the same command exited 0 without ncu).
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline14.txt; }
lap "full -m gpu suite, -x"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=5 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -12 $O/pytest_gpu_full.log
lap "default bench"
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_default.json') if l.startswith('{')][-1]); r=d['roofline']
print('value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f launches %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], r.get('launches'), d.get('parity')))
for kx in ('roofline_in_place_variant', 'roofline_streaming_variant', 'sustained'):
    print(kx, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in d.get(kx, {}).items() if a != 'what'})
print('clocks', d.get('clocks'))
PY
lap "launch list (2 steps) under ncu"
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/b_small.log 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_r02_shadow.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_launches.log 2>&1
echo "launch list exit $?"
lap "ncu --set full of the out-of-place pass"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_downdate_hyb -s 2 -c 1 -f -o $O/ncu_downdate_r02_shadow_rank32 \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_full.log 2>&1
echo "full capture exit $?"
ls -la $O/*.ncu-rep
lap done
