#!/bin/bash
# scan_ab.sh <workloads> <lib names...> — kernel_ms / roofline of the scan engine for each library variant ("-" = in-tree)
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
W=$1; shift
for lib in "$@"; do for w in $W; do
  if [ "$lib" = "-" ]; then unset RBACKNN_LIB; else export RBACKNN_LIB=$PWD/gpurun_tmp/librbacknn_$lib.so; fi
  python bench.py --workload $w --steps ${STEPS:-10} --warmup 3 --no-cpu --no-batched --no-single --no-legs 2>/dev/null | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('$lib $w', round(l['value']), round(l['roofline']['frac'],4), round(l['roofline']['kernel_ms'],4))"
done; done
