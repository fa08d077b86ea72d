#!/usr/bin/env bash
# r05 call 2 (1 GPU): GPU suite after the fixes, A/B of the short-item changes (producer lookahead, query rows
# through the item queue, deeper ring for 1-byte caches), cache-op lines.
set -uo pipefail
OUT=gpurun_out/r05b
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -12 $OUT/pytest_gpu.log
B="python bench.py --no-extra --no-cpu-baseline --steps 200 --warmup 5"
for wl in c4:v2 c4:v1 refbench:v1 refbench:v2 c3tp8:v1 c2:v1 c2:v2 small:v1 long:v2 c1:v1; do
  w=${wl%%:*}; v=${wl##*:}
  for cfg in new nolook noq stages16; do
    case $cfg in
      new) E="";;
      nolook) E="B200_ATTN_NO_LOOKAHEAD=1";;
      noq) E="B200_ATTN_NO_QBULK=1 B200_ATTN_NO_LOOKAHEAD=1";;
      stages16) E="B200_ATTN_MAX_STAGES=16";;
    esac
    env $E timeout 120 $B --workload $w --version $v > $OUT/bench_${w}_${v}_$cfg.json 2> $OUT/bench_${w}_${v}_$cfg.err
    python - "$OUT/bench_${w}_${v}_$cfg.json" "$w $v $cfg" <<'P'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d['roofline']
    print(sys.argv[2], 'value',d['value'],'attn_frac',r['frac'],'kernel_ms',r['kernel_ms'],'e2e',d['e2e']['value'])
except Exception as e: print(sys.argv[2],'FAILED',e)
P
  done
done
timeout 300 python tools/cache_ops_bench.py --iters 30 --only reshape,swap,convert > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err
echo "cache_ops exit $?"
python - <<'P'
import json
for l in open('gpurun_out/r05b/cache_ops.jsonl'):
    d=json.loads(l); print(d['case'], d.get('GBps'), d.get('frac_of_measured_hbm'), 'graph', d.get('graph_replay_GBps'), d.get('graph_replay_frac_of_measured_hbm'), d.get('bit_exact_vs_reference_cuda'))
P
