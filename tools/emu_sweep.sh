#!/bin/bash
# emu_sweep.sh — one GPU holding the stripe of rank 0 of 8 (C3): scan kernel time against the segment-cost constant
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
for c in ${COSTS:-10 20 40 80 160 320}; do
  RK_SEG_COST_X10=$c python bench.py --emu-world ${EMU:-8} --steps 20 --warmup 3 --no-cpu --no-batched --no-single --no-legs 2>/dev/null | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('cost_x10=$c', 'segs', l['run']['segments'], 'ms/step', round(l['ms_per_step'],4), 'scan_ms', round(l['roofline']['kernel_ms'],4), 'frac', round(l['roofline']['frac'],4))"
done
