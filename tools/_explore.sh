cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_parity.py -x -q 2>&1 | tail -3
for sm in 0 1; do
export RK_TENSOR_SAMPLE=$sm
echo "== sample $sm"
timeout 300 python tools/one_scan.py --rows 1000000 --dim 768 --kind cohere --metric ip --nq 1000 --filter none --engine tensor --reps 2 | tail -1
timeout 300 python tools/one_scan.py --rows 1000000 --dim 768 --kind cohere --metric ip --rbac random --nq 1000 --filter combination --engine tensor --reps 2 | tail -1
timeout 300 python tools/one_scan.py --rows 1000000 --nq 1000 --filter none --engine tensor --reps 2 | tail -1
timeout 300 python tools/one_scan.py --rows 10000000 --nq 1000 --filter none --engine tensor --reps 2| tail -1
timeout 300 python tools/one_scan.py --rows 10000000 --nq 1000 --filter bitmask --engine tensor --reps 2| tail -2
timeout 300 python tools/one_scan.py --rows 1000000 --nq 1000 --filter partition --engine tensor --reps 2| tail -2
done
