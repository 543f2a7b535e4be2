#!/bin/bash
# compute-sanitizer over every kernel variant.  Run on the GPU box:
#   gpurun --timeout 1500 -- 'bash scripts/gpu_sanitize.sh r02'
# Logs: gpurun_out/<tag>_sanitize_<tool>.log (copy the ones to keep into profiles/).
tag=${1:-r02}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" || exit 1
rc=0
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  timeout 900 compute-sanitizer --tool $tool $extra --error-exitcode 9 --print-limit 50 \
      python scripts/sanitize_cases.py ${SANITIZE_CASES} > gpurun_out/${tag}_sanitize_${tool}.log 2>&1
  e=$?
  echo "$tool exit $e: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${tag}_sanitize_${tool}.log | tail -1)"
  [ $e -ne 0 ] && rc=1
done
exit $rc
