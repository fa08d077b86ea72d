#!/usr/bin/env bash
# Repeat the whole GPU suite to catch intermittent failures (no -x: every failure is listed with its diagnosis).
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
for i in $(seq 1 ${REPS:-5}); do
  python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | grep -E "^(FAILED|E  +Assertion|[0-9]+ (passed|failed))|max err" | cut -c1-1200 > $OUT/flaky_$i.log
  echo "== run $i"; tail -8 $OUT/flaky_$i.log
done
