#!/usr/bin/env bash
# reshape_and_cache streamed variant: parity (cache tests, guard regions, reference-CUDA comparison) and A/B timing
set -uo pipefail
OUT=gpurun_out/r05rc
mkdir -p $OUT
timeout 300 python -m pytest tests/test_cache_gpu.py tests/test_guard_regions_gpu.py -k reshape_and_cache -m gpu -x -q > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/pytest.log
for cfg in "stream32x3:B200_X=1" "stream32x2:B200_RC_STREAM_STAGES=2" "stream16x3:B200_RC_STREAM_TOK=16" "nostream:B200_RC_NO_STREAM=1"; do
  name=${cfg%%:*}; envs=${cfg#*:}
  env $envs timeout 200 python tools/cache_ops_bench.py --iters 30 --only reshape > $OUT/rc_$name.jsonl 2> $OUT/rc_$name.err
  echo "$name exit $?"
  python - "$OUT/rc_$name.jsonl" <<'P'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l)
    if 'prefill' in d['case']: print('  ',d['case'],d['ms'],d['GBps'],d['frac_of_measured_hbm'],d.get('bit_exact_vs_reference_cuda'))
P
done
