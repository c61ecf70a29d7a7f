#!/bin/bash
# batched_ab.sh — the batched (RK_ENGINE_AUTO -> tensor) leg of c3 / c1 for the environment settings given in VARS ("A=1 B=2;A=0")
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
IFS=';' read -ra SETS <<< "${VARS:-RK_TENSOR_QTAU=1;RK_TENSOR_QTAU=0}"
for v in "${SETS[@]}"; do for w in ${WL:-c3 c1}; do
  env $v python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-single --no-legs 2>/dev/null | python -c "import sys,json; l=json.loads(sys.stdin.read()); b=l['batched']; print('$v $w batched', round(b['value']), 'ms/step', round(b['ms_per_step'],3), 'kernel', round(b['roofline']['kernel_ms'],3), 'frac', round(b['roofline']['frac'],4), 'same', b['identical_to_headline_results'], 'redo', b['tensor_redo'])"
done; done
