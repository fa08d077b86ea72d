#!/usr/bin/env bash
# ncu --set full of the attention kernel for the workloads in $WORKLOADS (default c4 c3tp8); each capture is
# preceded by the same command without ncu (profiling recipe).
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
for w in ${WORKLOADS:-c4 c3tp8}; do
  CMD="python bench.py --workload $w --steps 6 --warmup 3 --no-cpu-baseline"
  $CMD > $OUT/plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:paged_attn_ -s 4 -c 2 -f -o $OUT/prof_attn_$w $CMD > $OUT/ncu_full_$w.log 2>&1
  echo "ncu $w exit $?"
done
