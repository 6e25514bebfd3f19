#!/bin/bash
# ncu --set full captures of the two bandwidth kernels (run only after the same commands exited 0 without ncu):
#   gpurun --timeout 600 -- 'bash tools/gpu_ncu_full.sh'
# The .ncu-rep files come back in gpurun_out/; export them with
#   ncu -i X.ncu-rep --page details > profiles/X_details.txt ; ncu -i X.ncu-rep --page raw --csv > profiles/X_raw.csv
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_downdate -s 4 -c 2 -f -o $O/ncu_downdate_rank16 \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_full_10k.log 2>&1
echo "10k capture exit $?"
timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_mc_flush_w -s 3 -c 1 -f -o $O/ncu_mcflush \
    python bench.py --workload mc --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_full_mc.log 2>&1
echo "mc capture exit $?"
timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_mc_step -s 3 -c 1 -f -o $O/ncu_mcfront \
    python bench.py --workload mc --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_full_mcfront.log 2>&1
echo "mc front capture exit $?"
ls -la $O/*.ncu-rep
