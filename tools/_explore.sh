cd $GRAFT_REPO_ROOT
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c3.log 2> gpurun_out/bench_c3.err; tail -1 gpurun_out/bench_c3.log; tail -3 gpurun_out/bench_c3.err
python bench.py --workload c2 --steps 5 --warmup 3 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; tail -1 gpurun_out/bench_c2.log; tail -3 gpurun_out/bench_c2.err
python bench.py --workload c1 --steps 5 --warmup 3 > gpurun_out/bench_c1.log 2> gpurun_out/bench_c1.err; tail -1 gpurun_out/bench_c1.log; tail -3 gpurun_out/bench_c1.err
