#!/usr/bin/env bash
# ncu --set full on the 256-row prefill kernel (final tree), 2k + 2k shape
set -uo pipefail
OUT=gpurun_out/r05np
mkdir -p $OUT
timeout 60 python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 3 > $OUT/plain.log 2>&1; echo "plain exit $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'context_attn_tc256' -s 2 -c 1 -f -o $OUT/prof_tc256 \
    python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 3 > $OUT/ncu.log 2>&1
echo "ncu exit $?"; tail -2 $OUT/ncu.log | cut -c1-200
ncu -i $OUT/prof_tc256.ncu-rep --page raw --csv > $OUT/prof_tc256_raw.csv 2>/dev/null; wc -c $OUT/prof_tc256_raw.csv
