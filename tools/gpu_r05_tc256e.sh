#!/usr/bin/env bash
# inner timeouts sum to < 500 s
set -uo pipefail
OUT=gpurun_out/r05s
mkdir -p $OUT
for tn in 128 64; do
B200_CTX_TN=$tn timeout 90 python -m pytest tests/test_prefix_prefill_gpu.py tests/test_guard_regions_gpu.py -q --maxfail=10 -k "prefix or context" > $OUT/pytest_$tn.log 2>&1
echo "pytest tn=$tn exit $?"; tail -3 $OUT/pytest_$tn.log | cut -c1-250
B200_CTX_TN=$tn timeout 90 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill_tn$tn.jsonl 2>&1
echo "bench tn=$tn exit $?"
python - $tn <<'P'
import json,sys
for l in open(f'gpurun_out/r05s/prefix_prefill_tn{sys.argv[1]}.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print('tn',sys.argv[1], d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
done
CMD="python tools/prefix_prefill_bench.py llama7b_chunk2k_ctx2k_f16 2"
timeout 120 ncu --set full --clock-control none --import-source on -k regex:context_attn_tc256_kernel -s 2 -c 1 -f \
   -o $OUT/prof_ctx_tc256 $CMD > $OUT/ncu.log 2>&1
echo "ncu exit $?"
