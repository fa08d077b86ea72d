#!/usr/bin/env bash
set -uo pipefail
OUT=gpurun_out/r05k
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/pytest_gpu.log
B="python bench.py --no-extra --no-cpu-baseline --steps 200 --warmup 5"
for wl in c4:v2 c3tp8:v2 long:v2 c4:v1 c3tp8:v1 refbench:v1 small:v1; do
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
timeout 300 python tools/cache_ops_bench.py --iters 30 --only reshape,convert > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err
python - <<'P'
import json
for l in open('gpurun_out/r05k/cache_ops.jsonl'):
    d=json.loads(l); print(d['case'], d.get('GBps'), d.get('frac_of_measured_hbm'), 'graph', d.get('graph_replay_GBps'), d.get('graph_replay_frac_of_measured_hbm'), d.get('bit_exact_vs_reference_cuda'))
P
