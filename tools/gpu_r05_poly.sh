#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05t
mkdir -p $OUT
for poly in 1 2 0; do
B200_CTX_POLY=$poly timeout 90 python -m pytest tests/test_prefix_prefill_gpu.py -q --maxfail=10 > $OUT/pytest_poly$poly.log 2>&1
echo "pytest poly=$poly exit $?"; tail -3 $OUT/pytest_poly$poly.log | cut -c1-250
B200_CTX_POLY=$poly timeout 90 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_poly$poly.jsonl 2>&1
echo "bench poly=$poly exit $?"
python - $poly <<'P'
import json,sys
for l in open(f'gpurun_out/r05t/prefix_prefill_poly{sys.argv[1]}.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print('poly',sys.argv[1], d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
done
