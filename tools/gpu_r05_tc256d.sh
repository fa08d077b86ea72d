#!/usr/bin/env bash
# inner timeouts sum to < 600 s: a deadlocked kernel must not run the call into gpurun's limit
set -uo pipefail
OUT=gpurun_out/r05r
mkdir -p $OUT
timeout 100 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=10 -k "prefix or context" > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -8 $OUT/pytest.log | cut -c1-250
timeout 100 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1
echo "bench exit $?"
python - <<'P'
import json
for l in open('gpurun_out/r05r/prefix_prefill.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
CMD="python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 2"
timeout 150 ncu --set full --clock-control none --import-source on -k regex:context_attn_tc256_kernel -s 2 -c 1 -f \
   -o $OUT/prof_ctx_tc256 $CMD > $OUT/ncu.log 2>&1
echo "ncu exit $?"
