#!/bin/bash
# round_check.sh — one gpurun call: the GPU parity suite, smoke(), the default bench line and the single-query sweep.
# Everything lands in gpurun_out/check/.
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
O=gpurun_out/check; mkdir -p $O
( time timeout 1200 python -m pytest tests -x -q -m gpu ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
( time timeout 900 python bench.py ) > $O/bench_default.log 2>&1; echo "bench rc=$?" | tee -a $O/bench_default.log
timeout 300 python tools/single_query.py --reps 20 > $O/single_query_sweep.txt 2>&1; echo "single rc=$?"
tail -3 $O/pytest_gpu.log
