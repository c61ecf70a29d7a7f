set -u
R=r1
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 2>gpurun_out/c4.err > gpurun_out/c4.json; tail -2 gpurun_out/c4.err
python -c "
import json
d=json.loads(open('gpurun_out/c4.json').read()); r=d['roofline']; print(d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['probe']['ms'], r['list_scan_ms'], r['achieved'], d['config']['recall_at_100'], d['cpu_baseline']['value'], d['cpu_baseline']['gpu_ids_match'])"
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv --log-file gpurun_out/${R}_c4_launches.csv python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:rk_scan_kernel -s 3 -c 1 -o gpurun_out/${R}_scan_c4 -f python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:rk_ivf_keys -s 3 -c 1 -o gpurun_out/${R}_keys_c4 -f python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1
ls -la gpurun_out/${R}_*c4*
