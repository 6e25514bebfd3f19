#!/bin/bash
# Lean final check of the round: GPU tests, the two default bench lines, launch list and one full capture of the rank-32 flush.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/final; mkdir -p $O
timeout 400 python -m pytest tests/ -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "exit $?" >> $O/pytest_gpu.log; tail -2 $O/pytest_gpu.log
timeout 300 python bench.py > $O/bench_10k.json 2> $O/bench_10k.err
timeout 200 python bench.py --workload mc > $O/bench_mc.json 2> $O/bench_mc.err
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_10k.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/ncu_10k.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_downdate_tma -s 2 -c 1 -f -o $O/ncu_downdate_rank32 \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/ncu_full.log 2>&1
for f in bench_10k bench_mc; do python - $O/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(sys.argv[1], 'value %.5g e2e %.5g step %.4f K3 %.4f frac %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['avg_launch_ms'], r['frac']), d['clocks']['reasons'], d.get('roofline_streaming_variant'))
PY
done
