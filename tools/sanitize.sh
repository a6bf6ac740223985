#!/bin/bash
# compute-sanitizer over every kernel of libb2c.so (SURVEY section 5 "race detection"): one summary file per tool under
# $OUT (default gpurun_out/).  Usage: tools/sanitize.sh [tag] [tools...] [-- cases...]     (on a GPU box)
cd "$(dirname "$0")/.."
TAG=${1:-r2}; shift
TOOLS=(); CASES=()
while [ $# -gt 0 ] && [ "$1" != "--" ]; do TOOLS+=("$1"); shift; done
[ "$1" == "--" ] && shift
CASES=("$@")
[ ${#TOOLS[@]} -eq 0 ] && TOOLS=(memcheck racecheck synccheck initcheck)
[ ${#CASES[@]} -eq 0 ] && CASES=(lin_dare slots_tail slots_refill general rollout)
OUT=${OUT:-gpurun_out}
mkdir -p "$OUT"
CS=/usr/local/cuda/bin/compute-sanitizer
PY=$(python -c 'import sys, os; print(os.path.realpath(sys.executable))')
for tool in "${TOOLS[@]}"; do
  f="$OUT/sanitizer_${TAG}_${tool}.txt"
  : > "$f"
  for c in "${CASES[@]}"; do
    echo "=== $tool :: $c" >> "$f"
    raw="$OUT/.san_raw.txt"
    timeout 900 $CS --tool $tool --target-processes all --print-limit 20 --error-exitcode 0 "$PY" tools/sanitize_cases.py $c > "$raw" 2>&1
    rc=$?
    grep -E "case ok|ERROR SUMMARY|RACECHECK SUMMARY|========= (Error|Warning|Race|Invalid|Uninit|Barrier|Program hit|Hazard|Internal)|at .*\.cu:|AssertionError|Traceback" "$raw" | head -60 >> "$f"
    if ! grep -q "case ok" "$raw"; then echo "--- raw tail:" >> "$f"; tail -15 "$raw" >> "$f"; fi
    echo "--- exit $rc" >> "$f"
  done
done
rm -f "$OUT/.san_raw.txt"
grep -H -E "SUMMARY|case ok|exit" "$OUT"/sanitizer_${TAG}_*.txt | tail -80
