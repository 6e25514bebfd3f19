#!/bin/bash
# Round 2, GPU call 12 (1 GPU): continuous flush after the slot-release fix of k_downdate_hyb: full suite, A/B bench.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline13.txt; }
lap "full -m gpu suite, -x"
timeout 1500 python -m pytest tests/ -x -q -m gpu --durations=5 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -14 $O/pytest_gpu_full.log
lap "bench A/B"
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 > $O/bench_cf1.json 2> $O/bench_cf1.err
EKF_SHADOW=0 timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 100 --parity-steps 0 > $O/bench_cf0.json 2> $O/bench_cf0.err
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 20 > $O/bench_cf1_s20.json 2> $O/bench_cf1_s20.err
timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 200 --landmarks 1000 --parity-steps 0 > $O/bench_cf1_1k.json 2> $O/bench_cf1_1k.err
EKF_SHADOW=0 timeout 300 python bench.py --no-cpu-baseline --no-streaming-variant --sustained 0 --steps 200 --landmarks 1000 --parity-steps 0 > $O/bench_cf0_1k.json 2> $O/bench_cf0_1k.err
python - <<'PY'
import json
for f,what in [('bench_cf1','continuous flush'),('bench_cf0','EKF_SHADOW=0'),('bench_cf1_s20','continuous, 20 steps'),('bench_cf1_1k','1k continuous'),('bench_cf0_1k','1k EKF_SHADOW=0')]:
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1]); r=d['roofline']
        print(f, what, 'value %.0f e2e %.0f ms/step %.4f K3 %.4f frac %.3f launches %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'], r.get('launches'), d.get('parity')))
    except Exception as e: print(f, 'ERR', e, open('gpurun_out/%s.err'%f).read()[-800:])
PY
lap done
