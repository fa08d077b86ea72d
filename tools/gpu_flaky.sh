#!/usr/bin/env bash
# Repeat the attention parity tests to catch an intermittent failure (no -x: list every failure).
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
for i in 1 2 3 4; do
  python -m pytest tests/test_attention_gpu.py -q -m gpu -p no:cacheprovider 2>&1 | grep -v "^\s*$" | grep -E "^(FAILED|E  |[0-9]+ (passed|failed))|AssertionError" | cut -c1-900 > $OUT/flaky_$i.log
  echo "== run $i"; tail -12 $OUT/flaky_$i.log
done
python tools/cache_ops_bench.py > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err; echo "cache bench exit $?"
cut -c1-420 $OUT/cache_ops.jsonl; tail -3 $OUT/cache_ops.err
