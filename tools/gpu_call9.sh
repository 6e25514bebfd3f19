#!/bin/bash
# Round 2, GPU call 9 (1 GPU): full suite with the device-side measure; bench with lazy on/off, async measure on/off.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline9.txt; }
lap "device-side measure test first"
timeout 600 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "decided_on_device or lazy" > $O/pytest_am.log 2>&1; echo "exit $?" >> $O/pytest_am.log; tail -8 $O/pytest_am.log
lap "full -m gpu suite"
timeout 1500 python -m pytest tests/ -q -m gpu > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -8 $O/pytest_gpu_full.log
lap "bench variants"
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 > $O/bench_a.json 2> $O/bench_a.err
EKF_LAZY_FLUSH=0 timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 --parity-steps 0 > $O/bench_b.json 2> $O/bench_b.err
EKF_ASYNC_MEASURE=0 timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 --parity-steps 0 > $O/bench_c.json 2> $O/bench_c.err
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 200 --landmarks 1000 --parity-steps 0 > $O/bench_d.json 2> $O/bench_d.err
python - <<'PY'
import json
for f,what in [('bench_a','default'),('bench_b','lazy off'),('bench_c','async measure off'),('bench_d','1k landmarks')]:
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1]); r=d['roofline']
        print(f, what, 'value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], d.get('parity')))
    except Exception as e: print(f, 'ERR', e, open('gpurun_out/%s.err'%f).read()[-500:])
PY
lap done
