#!/bin/bash
# One gpurun call that re-measures the round: GPU parity tests, the two bench lines, the A/B switches of the
# decoupled state chain / early pose read-back / flush CTA width, and the ncu launch lists.  Every leg has its own
# timeout and writes under gpurun_out/; a failing leg does not stop the others.
#   gpurun --timeout 1100 -- 'bash tools/gpu_round_check.sh'
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline.txt; }

lap "pytest -m gpu"
timeout 780 python -m pytest tests/ -x -q -m gpu --durations=12 > $O/pytest_gpu.log 2>&1
echo "exit $?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
if ! grep -q "^exit 0" $O/pytest_gpu.log; then
    lap "pytest failed: same suite on the single-stream paths (EKF_ASYNC_X=0 EKF_MC_EARLY_POSES=0)"
    EKF_ASYNC_X=0 EKF_MC_EARLY_POSES=0 timeout 400 python -m pytest tests/test_parity_gpu.py tests/test_mc_gpu.py -x -q -m gpu \
        -k "not decoupled and not poses_readback and not tensor_core" > $O/pytest_gpu_sync.log 2>&1
    echo "exit $?" >> $O/pytest_gpu_sync.log
    tail -3 $O/pytest_gpu_sync.log
fi

lap "bench 10k (default)"
timeout 400 python bench.py > $O/bench_10k.json 2> $O/bench_10k.err
lap "bench 10k, EKF_ASYNC_X=0"
EKF_ASYNC_X=0 timeout 200 python bench.py --no-cpu-baseline --defer 8 > $O/bench_10k_sync.json 2> $O/bench_10k_sync.err
lap "bench mc (default)"
timeout 300 python bench.py --workload mc > $O/bench_mc.json 2> $O/bench_mc.err
lap "bench mc, EKF_MC_EARLY_POSES=0"
EKF_MC_EARLY_POSES=0 timeout 200 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_late.json 2> $O/bench_mc_late.err
lap "bench mc, default, 100 steps (for the A/B legs)"
timeout 200 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_100.json 2> $O/bench_mc_100.err
lap "bench mc, EKF_MC_TC=128"
EKF_MC_TC=128 timeout 200 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_tc128.json 2> $O/bench_mc_tc128.err
lap "bench mc, EKF_MC_TC=32"
EKF_MC_TC=32 timeout 200 python bench.py --workload mc --no-cpu-baseline --steps 100 > $O/bench_mc_tc32.json 2> $O/bench_mc_tc32.err

lap "10k other configurations (parity-test cases, short runs)"
EKF_GAIN_BLOCK=64 timeout 200 python bench.py --no-cpu-baseline --defer 8 > $O/bench_10k_gain64.json 2> $O/bench_10k_gain64.err
EKF_GAIN_BLOCK=32 timeout 200 python bench.py --no-cpu-baseline --defer 8 > $O/bench_10k_gain32.json 2> $O/bench_10k_gain32.err
timeout 200 python bench.py --no-cpu-baseline --defer 8 --storage full --steps 100 > $O/bench_10k_full.json 2> $O/bench_10k_full.err
timeout 200 python bench.py --no-cpu-baseline --defer 1 --steps 40 > $O/bench_10k_defer1.json 2> $O/bench_10k_defer1.err
timeout 200 python bench.py --no-cpu-baseline --landmarks 1000 > $O/bench_1k.json 2> $O/bench_1k.err
lap "ncu launch list 10k"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_10k.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_10k.log 2>&1
lap "ncu launch list mc"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_mc.csv \
    python bench.py --workload mc --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_mc.log 2>&1
lap "done"
for f in bench_10k bench_10k_sync bench_10k_gain64 bench_10k_gain32 bench_10k_full bench_10k_defer1 bench_1k bench_mc bench_mc_100 bench_mc_late bench_mc_tc128 bench_mc_tc32; do
    python - "$O/$f.json" <<'EOF'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d['roofline']
    print('%-28s value %.4g  e2e %.4g  ms/step %.4f  K3 %.4f ms  frac %.3f  clocks %s' % (
        sys.argv[1].split('/')[-1], d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac'],
        (d.get('clocks') or {}).get('reasons')))
except Exception as e:
    print(sys.argv[1], 'unreadable:', e)
EOF
done
