#!/bin/bash
# Round 2, GPU call 2 (1 GPU): the rewritten bench.py (both arms), launch list + ncu --set full of the hybrid rank-32 flush.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
lap() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a $O/timeline2.txt; }
lap "bench driver-like (20 steps)"
timeout 400 python bench.py --steps 20 --warmup 3 > $O/bench_10k_s20.json 2> $O/bench_10k_s20.err; tail -c 2500 $O/bench_10k_s20.json; tail -3 $O/bench_10k_s20.err
lap "bench default (200 steps), no cpu baseline"
timeout 400 python bench.py --no-cpu-baseline > $O/bench_10k_s200.json 2> $O/bench_10k_s200.err; tail -c 600 $O/bench_10k_s200.json
lap "reference arm"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; tail -c 1500 $O/bench_ref.json; tail -3 $O/bench_ref.err
lap "bench defer 8 / full"
timeout 200 python bench.py --no-cpu-baseline --defer 8 --steps 100 --sustained 0 --parity-steps 0 > $O/bench_10k_defer8.json 2> $O/bench_10k_defer8.err
timeout 200 python bench.py --no-cpu-baseline --storage full --steps 100 --sustained 0 > $O/bench_10k_full.json 2> $O/bench_10k_full.err
timeout 200 python bench.py --no-cpu-baseline --landmarks 1000 --steps 200 --sustained 0 > $O/bench_1k.json 2> $O/bench_1k.err
lap "launch list"
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/plain_ll.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_r02_defer16.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_ll.log 2>&1
lap "ncu --set full of the hybrid flush"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_downdate_hyb -s 2 -c 2 -f -o $O/ncu_downdate_r02_hyb_rank32 \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-streaming-variant --sustained 0 --parity-steps 0 > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
lap "bench mc"
timeout 300 python bench.py --workload mc --no-cpu-baseline > $O/bench_mc.json 2> $O/bench_mc.err; tail -c 400 $O/bench_mc.json
lap done
