#!/bin/bash
# compute-sanitizer targets for the CUDA kernels (run on a GPU box, one GPU):
#   tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck ...]   (default: memcheck racecheck synccheck)
# Every tool runs tools/sanitize_smoke.py (each kernel family once, emulated ranks). Reports go to
# gpurun_out/sanitize_<tool>.log; the exit code is non-zero if any tool reported an error.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export MP4J_GPU_TIMEOUT_S=${MP4J_GPU_TIMEOUT_S:-120}   # kernels run 10-100x slower under the sanitizer
tools=${@:-memcheck racecheck synccheck}
rc=0
for t in $tools; do
  timeout ${SANITIZE_TIMEOUT:-300} compute-sanitizer --tool $t --error-exitcode 9 python tools/sanitize_smoke.py \
    > gpurun_out/sanitize_$t.log 2>&1
  code=$?
  echo "$t: exit $code; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|sanitize_smoke ok' gpurun_out/sanitize_$t.log | tr '\n' ' ')"
  [ $code -ne 0 ] && rc=1
done
exit $rc
