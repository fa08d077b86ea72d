#!/usr/bin/env bash
# prefix prefill: grouped longest-first CTA order (ctx_cta_coords) -- parity, then the bench rows per group size
set -uo pipefail
OUT=gpurun_out/r05lpt
mkdir -p $OUT
#timeout 300 python -m pytest tests/test_prefix_prefill_gpu.py -q --maxfail=10 > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $OUT/pytest.log | cut -c1-250
for g in 64 128 4096; do
  B200_CTX_PAIR_GROUP=$g timeout 120 python tools/prefix_prefill_bench.py > $OUT/pp_g$g.jsonl 2>&1
  echo "group $g exit $?"
  python - $OUT/pp_g$g.jsonl <<'P'
import json,sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: continue
    print('   ', d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'])
P
done
