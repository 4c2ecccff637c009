#!/bin/bash
# compute-sanitizer over the small parity subset of scripts/sanitize_cases.py.
# memcheck: out-of-bounds / misaligned accesses; racecheck: shared-memory hazards of the y-sweep tile protocol;
# synccheck: divergent barriers; initcheck: reads of uninitialised device memory.  Logs -> gpurun_out/sanitize_*.log
mkdir -p gpurun_out
timeout 300 python scripts/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  timeout ${SAN_TIMEOUT:-1200} compute-sanitizer --tool $tool $extra --print-limit 40 python scripts/sanitize_cases.py "$@" \
      > gpurun_out/sanitize_$tool.log 2>&1
  echo "== $tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|SANITIZE_CASES_OK' gpurun_out/sanitize_$tool.log | tr '\n' ' ')"
done
