python -m pytest tests/test_gpu_subset.py tests/test_gpu_table.py -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -1
for r in 4 16 32; do
  echo "== replicas $r pair"; python tools/subset_bench.py --replicas $r --reps 4 | tail -1
done
