#!/usr/bin/env bash
# inner timeouts sum to < 600 s: a deadlocked kernel must not run the call into gpurun's limit
set -uo pipefail
OUT=gpurun_out/r05q
mkdir -p $OUT
timeout 150 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=10 -k "prefix or context" > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/pytest.log | cut -c1-250
for st in 1 0; do
B200_CTX_STAGGER=$st timeout 120 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_stagger$st.jsonl 2>&1
echo "bench stagger $st exit $?"
python - $st <<'P'
import json,sys
for l in open(f'gpurun_out/r05q/prefix_prefill_stagger{sys.argv[1]}.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print('stagger',sys.argv[1], d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
done
