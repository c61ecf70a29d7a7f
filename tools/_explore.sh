cd $GRAFT_REPO_ROOT
python bench.py --steps 3 --warmup 3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv --log-file gpurun_out/r1_bench_c3_launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_c3.log 2>&1
python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu > gpurun_out/plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:rk_ --csv --log-file gpurun_out/r1_bench_c2_launches.csv python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rk_bscan -s 4 -c 1 -o gpurun_out/r1_bscan_c2 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_c2f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rk_bscan -s 4 -c 1 -o gpurun_out/r1_bscan_c3 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_c3f.log 2>&1
ls -la gpurun_out | tail -12
