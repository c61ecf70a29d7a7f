#!/bin/bash
# capture_c3.sh — the C3 part of capture_profiles.sh alone (launch list + full captures of the batch scan and the one-CTA tensor kernel)
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
R=${ROUND:-r2}
OUT=gpurun_out/profiles_$R
TMP=/tmp/rk_ncu
mkdir -p $OUT $TMP
LIST="ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv"
FULL="ncu --set full --clock-control none --import-source on"
B="--no-legs --no-cpu"
python bench.py --steps 3 --warmup 3 $B > $TMP/plain_c3.log 2>&1 || { echo "bench c3 failed"; tail -5 $TMP/plain_c3.log; exit 1; }
grep '^{"metric"' $TMP/plain_c3.log > $OUT/plain_c3_line.json
timeout 600 $LIST -c 1500 --log-file $TMP/c3_launches.csv python bench.py --steps 3 --warmup 3 $B --single-reps 4 > /dev/null 2>&1
python tools/ncu_summary.py launches $TMP/c3_launches.csv $OUT/${R}_bench_c3_launches.txt
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
ls -la $OUT
