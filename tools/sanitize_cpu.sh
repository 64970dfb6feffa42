#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the host+device headers of the kernels (gpa_enum.cuh, part_table.cuh,
# subset_core.cuh, subset_plan.h), through the CPU harnesses of tests/csrc/ and their oracle-checked tests.
#   tools/sanitize_cpu.sh        (about 2 minutes; exits non-zero on the first report)
set -e
cd "$(dirname "$0")/.."
DIR=$(mktemp -d)
trap 'rm -rf "$DIR"' EXIT
export GPA_HARNESS_DIR="$DIR"
export GPA_HARNESS_CXXFLAGS="-O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined"
export LD_PRELOAD="$(g++ -print-file-name=libasan.so) $(g++ -print-file-name=libubsan.so)"
export ASAN_OPTIONS=detect_leaks=0:abort_on_error=1
python -m pytest tests/test_enum_core.py tests/test_subset_core.py -x -q
