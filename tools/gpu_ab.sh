#!/usr/bin/env bash
# A/B runs: each argument is "ENV=VAL,ENV2=VAL2:workload:version" (env part may be empty)
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
for spec in "$@"; do
  envs=${spec%%:*}; rest=${spec#*:}; w=${rest%%:*}; v=${rest##*:}
  ( IFS=','; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done
    python bench.py --workload $w --version $v --steps 200 --warmup 5 --no-cpu-baseline > $OUT/ab.json 2> $OUT/ab.err || tail -3 $OUT/ab.err
    python - <<PY
import json
d=json.load(open("$OUT/ab.json")); r=d["roofline"]
print("[$envs] $w $v kernel_ms", r["kernel_ms"], "GB/s", r["achieved"], "frac", r["frac"], "e2e", d["e2e"]["value"])
PY
  )
done
