#!/bin/bash
# compute-sanitizer over the golden fixtures on the GPU box: memcheck on the search + build + SBWT
# construction tests, racecheck on the build tests (the LCS rounds rely on "every destination byte has exactly one writer",
# include/lcs_basic_parallel_algorithm.hpp:76-78; the warp kernel stages edges in shared memory).
#   tools/sanitizer_run.sh gpurun_out/r2_sanitizer.log
out=${1:-gpurun_out/sanitizer.log}
mkdir -p "$(dirname "$out")"; : > "$out"
run() {
  echo "=== compute-sanitizer --tool $1 :: pytest $2 -k '$3'" >> "$out"
  timeout 900 compute-sanitizer --tool "$1" --target-processes all --error-exitcode 97 \
    python -m pytest $2 -m gpu -x -q -k "$3" > /tmp/san.log 2>&1
  echo "exit code $?" >> "$out"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Invalid|Race reported|hazard" /tmp/san.log | sort | uniq -c | head -20 >> "$out"
}
run memcheck tests/test_gpu_search.py "golden or known_answers"
run memcheck tests/test_gpu_build.py "golden_index_files or golden_lcs"
run memcheck tests/test_gpu_sbwt.py "golden"
run racecheck tests/test_gpu_build.py "golden_index_files or golden_lcs"
run racecheck tests/test_gpu_search.py "golden_search_output_as_runs"
cat "$out"
