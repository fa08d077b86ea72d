#!/usr/bin/env bash
# One gpurun call: GPU parity tests, cache-op bench lines, the default bench line.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/gpu.txt 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  ( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_gpu.log 2>&1
  echo "pytest exit $?"; tail -6 $OUT/pytest_gpu.log
fi
python tools/cache_ops_bench.py > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err; echo "cache bench exit $?"
cut -c1-420 $OUT/cache_ops.jsonl; tail -3 $OUT/cache_ops.err
python bench.py > $OUT/bench_c2_v1.json 2> $OUT/bench_c2_v1.err; echo "bench exit $?"; cat $OUT/bench_c2_v1.json
