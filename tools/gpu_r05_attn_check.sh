#!/usr/bin/env bash
# decode attention after a kernel change: the attention-side GPU tests, then the driver bench line with its extras
set -uo pipefail
OUT=gpurun_out/r05tp
mkdir -p $OUT
timeout 120 python -m pytest tests/test_attention_gpu.py -m gpu -x -q -k "gqa or mma or v2 or fp8" > $OUT/pytest_quick.log 2>&1
echo "quick exit $?"; tail -3 $OUT/pytest_quick.log
timeout 420 python -m pytest tests/test_attention_gpu.py tests/test_attention_long_gpu.py tests/test_guard_regions_gpu.py tests/test_decode_step_gpu.py tests/test_vs_reference_cuda_full_gpu.py tests/test_binding_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest.log 2>&1
echo "pytest exit $?"; tail -4 $OUT/pytest.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_n1.json 2> $OUT/bench_n1.err
echo "bench exit $?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r05tp/bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'roof',d['roofline']['frac'],'e2e',d['e2e']['value'],d['e2e']['frac_of_device_value'],'errors',d.get('extra_errors'))
for k,v in d.get('extra',{}).items(): print('  ',k,v['value'],v['roofline']['frac'],'e2e',v['e2e']['value'],v['e2e']['frac_of_device_value'])
P
