#!/bin/bash
# compute-sanitizer over the hand-synchronised kernels (SURVEY.md §5 "race detection"): racecheck (shared-memory
# hazards), synccheck (barrier misuse), memcheck (out-of-bounds / misaligned).  Run on the GPU box:
#     bash tools/sanitize.sh [outdir]            -> <outdir>/sanitizer_{racecheck,synccheck,memcheck}.txt
# Each tool gets its own process; the driver checks its results against the oracle.
out=${1:-gpurun_out}
mkdir -p "$out"
cd "$(dirname "$0")/.."
for tool in memcheck synccheck racecheck; do
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  timeout 900 compute-sanitizer --tool $tool $extra --print-limit 40 --kernel-name kns=3ibp \
      python tools/sanitize_driver.py "${@:2}" > "$out/sanitizer_$tool.txt" 2>&1
  echo "rc=$?" >> "$out/sanitizer_$tool.txt"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok:|rc=" "$out/sanitizer_$tool.txt" | sed "s/^/[$tool] /"
done
