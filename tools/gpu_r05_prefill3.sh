#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05f
mkdir -p $OUT
timeout 600 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=20 > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -8 $OUT/pytest.log
timeout 300 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1
python - <<'P'
import json
for l in open('gpurun_out/r05f/prefix_prefill.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['torch_sdpa_masked_ms'], d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
CMD="python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:context_attn_tc_kernel -s 2 -c 1 -f \
   -o $OUT/prof_ctx_tc $CMD > $OUT/ncu.log 2>&1
echo "ncu exit $?"
