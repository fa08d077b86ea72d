#!/usr/bin/env bash
# 256-bit stores in convert_fp8_e5m2 (fp8 -> 16-bit) and the streamed reshape_and_cache V path: parity + A/B
set -uo pipefail
OUT=gpurun_out/r05st
mkdir -p $OUT
timeout 300 python -m pytest tests/test_cache_gpu.py tests/test_guard_regions_gpu.py tests/test_rope_gpu.py tests/test_decode_step_gpu.py -k "reshape_and_cache or convert or rope or rotary or step" -m gpu -x -q > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $OUT/pytest.log
for cfg in "st256:B200_X=1"; do
  name=${cfg%%:*}; envs=${cfg#*:}
  env $envs timeout 200 python tools/cache_ops_bench.py --iters 30 --only reshape,convert,rope > $OUT/$name.jsonl 2> $OUT/$name.err
  echo "$name exit $?"
  python - "$OUT/$name.jsonl" <<'P'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l)
    if True: print('  ',d['case'],d['ms'],d['GBps'],d['frac_of_measured_hbm'],d.get('bit_exact_vs_reference_cuda'))
P
done
