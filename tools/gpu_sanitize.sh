#!/bin/bash
# compute-sanitizer passes over the native GPU tests that exercise the hand-written synchronisation: memcheck (out-of-
# bounds / misaligned global and shared accesses) and racecheck (shared-memory hazards: the TMA ring's mbarrier
# hand-over, the staging buffer of the copy kernel) on [exchange], [jacobi] and [astaroth]. Logs land in gpurun_out/sanitizer/;
# the summaries are copied to profiles/sanitizer/ by hand. The reference registers `cuda-memcheck test_cuda` as its
# CTest command (test/CMakeLists.txt:31,44); this is the same gate with today's tool.
mkdir -p gpurun_out/sanitizer
cd "$(dirname "$0")/.."
export STENCIL_LOG_LEVEL=1
# flag waits are bounded by wall time; under the sanitizer kernels run tens of times slower
export STENCIL_FLAG_TIMEOUT_MS=600000
S=gpurun_out/sanitizer/summary.txt
: > $S
for tool in memcheck racecheck; do
  for g in ${SAN_GROUPS:-"[pack]" "[exchange]" "[jacobi]" "[astaroth]"}; do
    name=$(echo "$g" | tr -d '[]')
    log=gpurun_out/sanitizer/${tool}_${name}.log
    timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 ./build/bin/test_cuda "$g" > $log 2>&1
    rc=$?
    echo "$tool $g rc=$rc | $(grep -E "test cases" $log | tail -1) | $(grep -E "ERROR SUMMARY|RACECHECK SUMMARY" $log | tail -1)" >> $S
  done
done
cat $S
