#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05p
mkdir -p $OUT
timeout 600 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=10 -k "prefix or context" > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/pytest.log | cut -c1-250
timeout 300 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1
python - <<'P'
import json
for l in open('gpurun_out/r05p/prefix_prefill.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['torch_sdpa_masked_ms'], d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
timeout 200 python bench.py --workload n1 > $OUT/bench_n1.json 2> $OUT/bench_n1.err; cat $OUT/bench_n1.json | cut -c1-900
