#!/usr/bin/env bash
# prefix prefill after the loader changes (tiles announced ahead of the refill, block numbers fetched before the wait)
set -uo pipefail
OUT=gpurun_out/r05ld
mkdir -p $OUT
timeout 400 python -m pytest tests/test_prefix_prefill_gpu.py -q -x > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -2 $OUT/pytest.log | cut -c1-200
timeout 120 python tools/prefix_prefill_bench.py > $OUT/pp_default.jsonl 2>&1; echo "bench exit $?"
B200_CTX_ATTN=tc128 timeout 120 python tools/prefix_prefill_bench.py > $OUT/pp_tc128.jsonl 2>&1; echo "tc128 exit $?"
python - <<'P'
import json
for f in ('pp_default','pp_tc128'):
    for l in open(f'gpurun_out/r05ld/{f}.jsonl'):
        try: d=json.loads(l)
        except Exception: continue
        print(f, d['case'], d['ms'], d['TFLOPs'], d['frac_of_measured_bf16_peak'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
