#!/usr/bin/env bash
# r05: ncu of the decode kernel under the v2 signature (quad mode) vs v1 on C4, and the launch list of the bench command
set -uo pipefail
OUT=gpurun_out/r05h
mkdir -p $OUT
for v in v2 v1; do
  CMD="python tools/attn_once.py --workload c4 --version $v --iters 4"
  timeout 120 $CMD > $OUT/plain_$v.log 2>&1; echo "plain $v exit $?"
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:'paged_attn' -s 2 -c 3 -f -o $OUT/prof_c4_$v $CMD > $OUT/ncu_$v.log 2>&1
  echo "ncu $v exit $?"
done
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-extra"
timeout 200 $CMD > $OUT/plain_bench.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "launch list exit $?"
