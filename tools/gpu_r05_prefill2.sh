#!/usr/bin/env bash
# r05: prefix prefill with a 3-stage K/V ring; full GPU suite after the planner / quad-threshold changes.
set -uo pipefail
OUT=gpurun_out/r05e
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -6 $OUT/pytest_gpu.log
timeout 300 python tools/prefix_prefill_bench.py > $OUT/prefix_prefill.jsonl 2>&1
python - <<'P'
import json
for l in open('gpurun_out/r05e/prefix_prefill.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d['case'], d['ms'], d['TFLOPs'], 'sdpa', d['torch_sdpa_masked_ms'], d['speedup_vs_torch_sdpa'], d['max_abs_diff_vs_sdpa'])
P
B="python bench.py --no-extra --no-cpu-baseline --steps 200 --warmup 5"
for wl in refbench:v1 small:v1 refbench:v2 small:v2 c4:v2 c3tp8:v2 c1:v1; do
  w=${wl%%:*}; v=${wl##*:}
  timeout 120 $B --workload $w --version $v > $OUT/bench_${w}_${v}.json 2> $OUT/bench_${w}_${v}.err
  python - "$OUT/bench_${w}_${v}.json" "$w $v" <<'P'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d['roofline']
    print(sys.argv[2], 'value',d['value'],'attn_frac',r['frac'],'kernel_ms',r['kernel_ms'],'e2e',d['e2e']['value'])
except Exception as e: print(sys.argv[2],'FAILED',e)
P
done
