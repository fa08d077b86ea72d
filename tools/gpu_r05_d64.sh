#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05v
mkdir -p $OUT
timeout 300 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=10 -k "prefix or context or 256 or variants or golden or head" > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -12 $OUT/pytest.log | cut -c1-250
timeout 120 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1
echo "bench exit $?"
python - <<'P'
import json
for l in open('gpurun_out/r05v/prefix_prefill.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
