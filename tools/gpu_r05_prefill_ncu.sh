#!/usr/bin/env bash
# r05: source-level ncu capture of the tcgen05 prefix-prefill kernel (where do the softmax warps wait?)
set -uo pipefail
OUT=gpurun_out/r05d
mkdir -p $OUT
CMD="python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 2"
timeout 200 $CMD > $OUT/plain.log 2>&1; echo "plain exit $?"; cat $OUT/plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:context_attn_tc_kernel -s 2 -c 1 -f \
   -o $OUT/prof_ctx_tc $CMD > $OUT/ncu.log 2>&1
echo "ncu exit $?"; tail -2 $OUT/ncu.log
timeout 300 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1; cat $OUT/prefix_prefill.jsonl | cut -c1-330
