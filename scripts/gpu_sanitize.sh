#!/bin/bash
# compute-sanitizer over the hand-rolled mbarrier / TMEM / cluster kernels on the smallest end-to-end cases.
# ONE tool per gpurun call (B200_PROFILING.md): usage  scripts/gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck [tag]
# The summary (last lines of the log + the ERROR SUMMARY line) goes to gpurun_out/<tag>_sanitize_<tool>.txt; copy it
# to profiles/ to keep it.
TOOL=${1:-memcheck}
TAG=${2:-r3}
O=gpurun_out
mkdir -p $O
SEL='tests/test_gpu_path.py::test_path_tiny64 tests/test_gpu_path.py::test_path_tiny80 tests/test_gpu_wire.py::test_segment_bits_matches_segment_and_fused_metrics'
# the plain run must pass first (a faulting program under a tool can take the GPU down)
timeout 600 python -m pytest $SEL -x -q > $O/${TAG}_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/${TAG}_sanitize_plain.log; exit 1; }
timeout 700 compute-sanitizer --tool $TOOL --launch-timeout 300 --error-exitcode 3 \
    python -m pytest $SEL -x -q -p no:cacheprovider > $O/${TAG}_sanitize_${TOOL}.log 2>&1
RC=$?
{ echo "# compute-sanitizer --tool $TOOL  rc=$RC  ($(date -u +%Y-%m-%dT%H:%MZ))  tests: $SEL";
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" $O/${TAG}_sanitize_${TOOL}.log | tail -20; } > $O/${TAG}_sanitize_${TOOL}.txt
cat $O/${TAG}_sanitize_${TOOL}.txt
exit 0
