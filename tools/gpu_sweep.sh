#!/usr/bin/env bash
# Sweep the v1 partition count (B200_ATTN_SPLIT) per workload; prints kernel ms / GB/s.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
for w in ${WORKLOADS:-c3tp8 c3tp2 c4}; do
  for np in ${SPLITS:-0 1 2 3 4 6 8 12 16}; do
    if [ "$np" = "0" ]; then unset B200_ATTN_SPLIT; else export B200_ATTN_SPLIT=$np; fi
    python bench.py --workload $w --steps 200 --warmup 5 --no-cpu-baseline > $OUT/sweep.json 2> $OUT/sweep.err || { tail -3 $OUT/sweep.err; continue; }
    python - <<PY
import json
d=json.load(open("$OUT/sweep.json")); r=d["roofline"]
print("$w np=$np kernel_ms", r["kernel_ms"], "GB/s", r["achieved"], "frac", r["frac"])
PY
  done
done
