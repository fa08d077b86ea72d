#!/usr/bin/env bash
# ncu --set full on the streamed reshape_and_cache kernel and the 256-bit-store convert kernel (one GPU)
set -uo pipefail
OUT=gpurun_out/r05nc
mkdir -p $OUT
timeout 100 python tools/cache_ops_bench.py --iters 1 --ncu --only reshape,convert > $OUT/plain.log 2>&1; echo "plain exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'reshape_and_cache_stream|from_fp8' \
    -c 12 -f -o $OUT/prof_cache2 python tools/cache_ops_bench.py --iters 1 --ncu --only reshape,convert > $OUT/ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/ncu.log
ncu -i $OUT/prof_cache2.ncu-rep --page raw --csv > $OUT/prof_cache2_raw.csv 2>/dev/null; wc -c $OUT/prof_cache2_raw.csv
