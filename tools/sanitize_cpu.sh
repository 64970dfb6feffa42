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

# Second leg: the library's HOST code (C ABI wrappers, replica.cpp, plugin.cpp, nvml_host.cpp) built with the same
# sanitizers, under the host / wire / lifecycle tests (none of them needs a GPU).
LIBDIR=$(mktemp -d)
trap 'rm -rf "$DIR" "$LIBDIR"' EXIT
(cd k8s-gpu-sharing-plugin_b200/csrc && env -u LD_PRELOAD /usr/local/cuda/bin/nvcc -O1 -g -std=c++17 -gencode arch=compute_100a,code=sm_100a \
  -Xcompiler -fPIC,-fsanitize=address,-fsanitize=undefined,-fno-sanitize-recover=undefined -shared -o "$LIBDIR/libgpualloc.so" \
  gpualloc.cu probes.cu replica.cpp plugin.cpp nvml_host.cpp -lpthread -ldl)
export GPA_TEST_LIB="$LIBDIR/libgpualloc.so"
export ASAN_OPTIONS=detect_leaks=0:abort_on_error=1:protect_shadow_gap=0
python -m pytest tests/test_abi_host.py tests/test_host_plugin_nvml.py tests/test_lifecycle.py tests/test_wire_grpc.py -x -q
