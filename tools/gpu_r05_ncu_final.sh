#!/usr/bin/env bash
# r05 final ncu evidence (1 GPU): launch list of the driver's bench command, full captures of the decode kernels
# (C2 SIMT, C3 tp=8 MMA, C4 v1 and v2 = quad mode + reduce), cache-op lines.
set -uo pipefail
OUT=gpurun_out/r05u
mkdir -p $OUT
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-extra"
timeout 100 $CMD > $OUT/plain_bench.log 2>&1 && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "launch list exit $?"
for spec in "c2 v1" "c3tp8 v1" "c4 v1" "c4 v2"; do
  set -- $spec
  C="python tools/attn_once.py --workload $1 --version $2 --iters 4"
  timeout 60 $C > $OUT/plain_$1_$2.log 2>&1; echo "plain $1 $2 exit $?"
  timeout 150 ncu --set full --clock-control none --import-source on -k regex:'paged_attn' -s 2 -c 2 -f -o $OUT/prof_$1_$2 $C > $OUT/ncu_$1_$2.log 2>&1
  echo "ncu $1 $2 exit $?"
done
timeout 200 python tools/cache_ops_bench.py --iters 30 --only reshape,convert,swap,copy > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err
python - <<'P'
import json
for l in open('gpurun_out/r05u/cache_ops.jsonl'):
    d=json.loads(l); print(d['case'], d.get('GBps'), d.get('frac_of_measured_hbm'), 'graph', d.get('graph_replay_GBps'), d.get('graph_replay_frac_of_measured_hbm'), d.get('bit_exact_vs_reference_cuda'))
P
