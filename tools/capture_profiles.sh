#!/bin/bash
# capture_profiles.sh — the ncu passes behind profiles/ (run on the GPU box: gpurun -- 'bash tools/capture_profiles.sh').
# Each pass only after the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
# Summaries: tools/ncu_summary.py (launch lists, key metrics), tools/ncu_stalls.py, tools/ncu_lines.py.
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
R=${ROUND:-r1}
LIST="ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv"
FULL="ncu --set full --clock-control none --import-source on"
for w in c3 c2 c5 c4; do
    python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/plain_$w.log 2>&1 || { echo "bench $w failed"; exit 1; }
    timeout 400 $LIST --log-file gpurun_out/${R}_${w}_launches.csv python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1
done
# dominant kernels, one launch each (-s skips the warm-up launches)
timeout 400 $FULL -k regex:rk_scan_kernel -s 4 -c 1 -o gpurun_out/${R}_scan_c3 -f python bench.py --steps 2 --warmup 3 --no-cpu --no-batched > /dev/null 2>&1
timeout 400 $FULL -k regex:rk_bscan2 -s 5 -c 1 -o gpurun_out/${R}_bscan2_c2 -f python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1
timeout 400 $FULL -k regex:rk_bscan_kernel -s 4 -c 1 -o gpurun_out/${R}_bscan_c3 -f python bench.py --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1
timeout 300 $FULL -k regex:rk_scan_kernel -s 2 -c 1 -o gpurun_out/${R}_scan_single -f python tools/one_scan.py --rows 10000000 --nq 1 --filter none --reps 4 > /dev/null 2>&1
timeout 500 $FULL -k regex:rk_scan_kernel -s 3 -c 1 -o gpurun_out/${R}_scan_c4 -f python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1
timeout 500 $FULL -k regex:rk_ivf_keys -s 3 -c 1 -o gpurun_out/${R}_keys_c4 -f python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1
ls -la gpurun_out/${R}_*
