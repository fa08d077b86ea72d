#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05w
mkdir -p $OUT
timeout 300 python -m pytest tests/test_prefix_prefill_gpu.py -q --maxfail=10 > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -4 $OUT/pytest.log | cut -c1-250
B200_CTX_ATTN=tc128 timeout 120 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_tc128.jsonl 2>&1
echo "bench exit $?"
python - <<'P'
import json
for l in open('gpurun_out/r05w/prefix_prefill_tc128.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print('tc128', d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
