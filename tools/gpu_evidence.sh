#!/usr/bin/env bash
# One gpurun call for a round's single-GPU evidence: parity tests, every bench line, cache-op lines,
# ncu launch list + full captures (C2 SIMT, C4 / C3tp8 MMA).  Results in gpurun_out/.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
SKIP_NCU=1 bash tools/gpu_check.sh
python tools/cache_ops_bench.py > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err; echo "cache bench exit $?"
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
for w in c2 c4 c3tp8; do
  CMD="python bench.py --workload $w --steps 6 --warmup 3 --no-cpu-baseline"
  $CMD > $OUT/plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:paged_attn_kernel -s 4 -c 2 -f -o $OUT/prof_attn_$w $CMD > $OUT/ncu_full_$w.log 2>&1
  echo "ncu full $w exit $?"
done
