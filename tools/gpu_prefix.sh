#!/usr/bin/env bash
# One gpurun call: prefix-prefill parity tests + throughput lines.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
( time timeout 900 python -m pytest tests/test_prefix_prefill_gpu.py -x -q ) > $OUT/pytest_prefix.log 2>&1
echo "pytest exit $?"; tail -15 $OUT/pytest_prefix.log
timeout 600 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2> $OUT/prefix_prefill.err; echo "bench exit $?"
cat $OUT/prefix_prefill.jsonl; tail -3 $OUT/prefix_prefill.err
