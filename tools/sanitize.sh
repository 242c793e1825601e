#!/bin/bash
# compute-sanitizer pass over the stand-alone kernel self-test (the reference has no
# sanitizer runs; the hand-written kernels here get memcheck / synccheck / initcheck / racecheck).
# Usage (on a GPU box): tools/sanitize.sh [out-file]      - needs `make -C csrc` first.
# NOTE: on the shared B200 pool of this build compute-sanitizer is administratively closed (the
# gpurun wrapper refuses it), so no summary is committed under profiles/; the script is for boxes
# where the tool may run. Each case is bounded by `timeout`; a tool that cannot instrument a kernel is reported, not fatal.
out=${1:-gpurun_out/sanitizer_summary.txt}
mkdir -p "$(dirname "$out")"
: > "$out"
SAN=${SANITIZER:-compute-sanitizer}
run() {   # tool, case
  local log; log=$(mktemp)
  timeout 300 "$SAN" --tool "$1" --error-exitcode 9 ./build/selftest "$2" > "$log" 2>&1
  local rc=$?
  local summary; summary=$(grep -E "ERROR SUMMARY|RACECHECK SUMMARY" "$log" | tail -1)
  local verdict; verdict=$(grep -E "^selftest:" "$log" | tail -1)
  echo "[$1] $2: exit=$rc | ${summary:-no summary line} | ${verdict:-no selftest verdict}" >> "$out"
  if [ $rc -ne 0 ]; then grep -E "=========" "$log" | head -20 >> "$out"; fi
  rm -f "$log"
}
for c in gemm1x1 conv3x3_a conv3x3_partial conv1x7 conv3x3_c20 gemm1x1_k40 conv3x3_k72 \
         conv3x3_splitk_in conv3x3_splitk_bn linear_2048_1000; do
  run memcheck "$c"
done
for c in conv3x3_a conv3x3_splitk_bn; do
  run synccheck "$c"
  run initcheck "$c"
done
run racecheck conv3x3_a
cat "$out"
