#!/bin/bash
# capture_profiles.sh — the ncu passes behind profiles/ (run on the GPU box: gpurun -- 'bash tools/capture_profiles.sh').
# Each pass only after the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
# The raw reports stay in /tmp on the box (together they exceed what gpurun copies back); their summaries — launch lists
# (tools/ncu_summary.py launches), key metrics + stall mix (ncu_summary.py full), stall samples per source line
# (ncu_lines.py) and per SASS instruction (ncu_stalls.py) — are written to gpurun_out/profiles_<round>/.
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
R=${ROUND:-r2}
OUT=gpurun_out/profiles_$R
TMP=/tmp/rk_ncu
mkdir -p $OUT $TMP
LIST="ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv"
FULL="ncu --set full --clock-control none --import-source on"
B="--no-legs --no-cpu"
# launch lists: the default line's headline + batched + single-query legs, then each other workload on its own
python bench.py --steps 3 --warmup 3 $B > $TMP/plain_c3.log 2>&1 || { echo "bench c3 failed"; tail -5 $TMP/plain_c3.log; exit 1; }
timeout 600 $LIST -c 1500 --log-file $TMP/c3_launches.csv python bench.py --steps 3 --warmup 3 $B --single-reps 4 > /dev/null 2>&1
python tools/ncu_summary.py launches $TMP/c3_launches.csv $OUT/${R}_bench_c3_launches.txt
for w in c1 c2 c5 c4; do
    python bench.py --workload $w --steps 3 --warmup 3 $B > $TMP/plain_$w.log 2>&1 || { echo "bench $w failed"; exit 1; }
    timeout 400 $LIST --log-file $TMP/${w}_launches.csv python bench.py --workload $w --steps 3 --warmup 3 $B > /dev/null 2>&1
    python tools/ncu_summary.py launches $TMP/${w}_launches.csv $OUT/${R}_bench_${w}_launches.txt
done
full() {   # name, kernel regex, skip, command...
    local name=$1 kern=$2 skip=$3; shift 3
    timeout 500 $FULL -k regex:$kern -s $skip -c 1 -o $TMP/$name -f "$@" > /dev/null 2>&1
    if [ -f $TMP/$name.ncu-rep ]; then
        python tools/ncu_summary.py full $TMP/$name.ncu-rep $OUT/${R}_${name}_ncu_full_metrics.json
        python tools/ncu_lines.py $TMP/$name.ncu-rep 40 > $OUT/${R}_${name}_lines.txt 2>&1
        python tools/ncu_stalls.py $TMP/$name.ncu-rep 30 > $OUT/${R}_${name}_stalls.txt 2>&1
    else echo "capture $name failed"; fi
}
S="--no-single"
full scan_c3 rk_scan_kernel 4 python bench.py --steps 2 --warmup 3 $B $S --no-batched
full bscan_c3 rk_bscan_kernel 4 python bench.py --steps 3 --warmup 3 $B $S
full bscan2_c2 rk_bscan2 5 python bench.py --workload c2 --steps 3 --warmup 3 $B
full scan_c4 rk_scan_kernel 3 python bench.py --workload c4 --steps 2 --warmup 3 $B
# the fused single-query kernel (tools/single_query.py: 2 + 2 x 5 warm-up calls first): an unfiltered full-corpus query,
# then a filtered query of role depth 1
full single_full rk_scan_single 6 python tools/single_query.py --reps 2
full single_depth1 rk_scan_single 40 python tools/single_query.py --reps 2
ls -la $OUT
