#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05j
mkdir -p $OUT
timeout 900 python -m pytest tests/test_attention_gpu.py tests/test_attention_long_gpu.py tests/test_guard_regions_gpu.py tests/test_decode_step_gpu.py tests/test_vs_reference_cuda_full_gpu.py -q --maxfail=20 > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -4 $OUT/pytest.log
B="python bench.py --no-extra --no-cpu-baseline --steps 200 --warmup 5"
for wl in c4:v2 c3tp8:v2 long:v2 c3:v2 c2:v2 c4:v1; do
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
