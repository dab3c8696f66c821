#!/bin/bash
# compute-sanitizer passes over one small shape of every kernel family (run under gpurun); logs -> gpurun_out/.
#   tools/sanitize.sh [tool ...]      default: memcheck racecheck synccheck initcheck
set -u
mkdir -p gpurun_out
TOOLS="${*:-memcheck racecheck synccheck initcheck}"
python tools/sanitize_run.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
for t in $TOOLS; do
  extra=""
  [ "$t" = "memcheck" ] && extra="--leak-check no"
  # only this library's kernels: torch's own fill/copy kernels are not under test
  timeout 1500 compute-sanitizer --tool "$t" $extra --kernel-regex kns=llie --print-limit 30 \
      python tools/sanitize_run.py > "gpurun_out/sanitize_$t.log" 2>&1
  echo "== $t rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_$t.log | tail -1)"
  grep -E "all families done" "gpurun_out/sanitize_$t.log" || echo "   (run did not finish)"
done
