python -m pytest tests/test_gpu_subset.py -m gpu -x -q 2>&1 | tail -3
for r in 4 16 32; do
  echo "== replicas $r pair"; python tools/subset_bench.py --replicas $r --reps 4 | tail -1
done
