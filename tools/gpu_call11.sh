#!/bin/bash
# Round 2, GPU call 11 (1 GPU): continuous flush (out-of-place K3 pass on its own stream + robot strip): tests, then A/B bench.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline11.txt; }
lap "continuous-flush tests"
timeout 600 python -m pytest tests/test_round2_gpu.py -x -q -m gpu -k "continuous or lazy or decided_on_device or failed_measure or nonfinite or writable" > $O/pytest_cf.log 2>&1; echo "exit $?" >> $O/pytest_cf.log; tail -25 $O/pytest_cf.log
lap "defer-16 parity at 1k x 1000 steps and 10k x 100 steps"
timeout 900 python -m pytest tests/test_round2_gpu.py tests/test_parity_gpu.py -x -q -m gpu -k "config2 or config3" > $O/pytest_cf2.log 2>&1; echo "exit $?" >> $O/pytest_cf2.log; tail -12 $O/pytest_cf2.log
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
