#!/usr/bin/env bash
# The no-allocation v2 entry point: its test alone in a fresh process, then the whole GPU suite the way the driver
# runs it (-x; tests/test_zz_no_alloc.py is collected last), smoke() and the driver's bench command.
#   gpurun --timeout 1200 -- 'bash tools/gpu_r05_noalloc.sh'
set -uo pipefail
OUT=gpurun_out/r05noalloc
mkdir -p $OUT
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_alone.log 2>&1
echo "alone exit $?"; tail -15 $OUT/pytest_alone.log
timeout 100 python -m pytest tests -m gpu --collect-only -q 2>/dev/null | tail -4
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1
echo "suite exit $?"; tail -6 $OUT/pytest_gpu.log
timeout 100 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_n1.json 2> $OUT/bench_n1.err
echo "bench exit $?"; cut -c1-400 $OUT/bench_n1.json
