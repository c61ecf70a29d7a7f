cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/one_scan.py --rows 1000000 --dim 768 --kind cohere --metric ip --rbac random --nq 1000 --filter combination --engine tensor --reps 2 | tail -1
timeout 300 python tools/one_scan.py --rows 10000000 --nq 1000 --filter bitmask --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 1000000 --nq 1000 --filter partition --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 100 --filter bitmask --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 100 --filter bitmask --engine scan --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 16 --filter bitmask --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 16 --filter bitmask --engine scan --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 1 --filter bitmask --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 10000000 --nq 1 --filter bitmask --engine scan --reps 2| tail -2
