#!/usr/bin/env bash
# Quick GPU iteration on the attention kernels: attention tests, then GQA bench lines.
set -uo pipefail
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_attention_gpu.py -x -q > $OUT/pytest_attn.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest_attn.log
for wl in ${WORKLOADS:-c3tp8:v1 c3tp2:v1 c3:v1 c4:v1 c4:v2 c2:v1}; do
  w=${wl%%:*}; v=${wl##*:}
  python bench.py --workload $w --version $v --steps 200 --warmup 5 --no-cpu-baseline > $OUT/bench_${w}_${v}.json 2> $OUT/bench_${w}_${v}.err
  echo "bench $w $v exit $?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${w}_${v}.json")); r=d["roofline"]
    print("  ", d["config"]["workload"][:40], "kernel_ms", r["kernel_ms"], "GB/s", r["achieved"], "frac", r["frac"], "e2e", d["e2e"]["value"], "clk", d["clocks"]["sm_mhz"])
except Exception as e:
    print("  no json", e); print(open("$OUT/bench_${w}_${v}.err").read()[-1500:])
PY
done
