#!/usr/bin/env bash
# r05 call 4 (1 GPU): quad mode (one v2 partition per consumer warp) parity + A/B, v1 split sweep for tiny batches.
set -uo pipefail
OUT=gpurun_out/r05c
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -12 $OUT/pytest_gpu.log
B="python bench.py --no-extra --no-cpu-baseline --steps 200 --warmup 5"
run() {  # label, env..., -- workload version
  local label=$1; shift; local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 120 $B --workload $1 --version $2 > $OUT/bench_$1_$2_$label.json 2> $OUT/bench_$1_$2_$label.err
  python - "$OUT/bench_$1_$2_$label.json" "$1 $2 $label" <<'P'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d['roofline']
    print(sys.argv[2], 'value',d['value'],'attn_frac',r['frac'],'kernel_ms',r['kernel_ms'],'e2e',d['e2e']['value'])
except Exception as e: print(sys.argv[2],'FAILED',e)
P
}
for wl in c4:v2 refbench:v2 long:v2 c3tp8:v2 c3:v2 small:v2; do
  w=${wl%%:*}; v=${wl##*:}
  run quad X=1 -- $w $v
  run noquad B200_ATTN_NO_QUAD=1 -- $w $v
  run quad3cta B200_ATTN_CTAS_PER_SM=3 -- $w $v
done
for np in 1 4 6 8 10 12 16 24; do
  run split$np B200_ATTN_SPLIT=$np -- refbench v1
  run split$np B200_ATTN_SPLIT=$np -- small v1
done
run default X=1 -- refbench v1
run default X=1 -- small v1
run default X=1 -- c4 v1
