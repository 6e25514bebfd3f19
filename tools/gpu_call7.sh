#!/bin/bash
# Round 2, GPU call 7 (1 GPU): full -m gpu suite; Monte-Carlo front with / without the robot-row cache.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline7.txt; }
lap "full -m gpu suite"
timeout 1500 python -m pytest tests/ -q -m gpu --durations=8 > $O/pytest_gpu_full.log 2>&1; echo "exit $?" >> $O/pytest_gpu_full.log; tail -16 $O/pytest_gpu_full.log
lap "bench mc: robot cache on / off"
timeout 300 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_rc1.json 2> $O/bench_mc_rc1.err
EKF_MC_ROBOT_CACHE=0 timeout 300 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_rc0.json 2> $O/bench_mc_rc0.err
python - <<'PY'
import json
for f in ['bench_mc_rc1','bench_mc_rc0']:
    try:
        d=json.loads([l for l in open('gpurun_out/%s.json'%f) if l.startswith('{')][-1]); r=d['roofline']
        print(f, 'value %.4g e2e %.4g ms/step %.4f flush %.4f front %.4f frac %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['front_kernel_avg_ms'], r['frac']))
    except Exception as e: print(f, 'ERR', e, open('gpurun_out/%s.err'%f).read()[-600:])
PY
lap done
