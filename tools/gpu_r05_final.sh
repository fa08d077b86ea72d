#!/usr/bin/env bash
# r05 final single-GPU pass: whole GPU suite, smoke(), the driver's bench command, the prefix-prefill bench row
set -uo pipefail
OUT=gpurun_out/r05final
mkdir -p $OUT
timeout 500 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -4 $OUT/pytest_gpu.log
timeout 100 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_n1.json 2> $OUT/bench_n1.err
echo "bench exit $?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r05final/bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'roof',d['roofline']['frac'],'e2e',d['e2e']['value'],d['e2e']['frac_of_device_value'],'errors',d.get('extra_errors'))
for k,v in d.get('extra',{}).items(): print('  ',k,v['value'],v['roofline']['frac'],'e2e',v['e2e']['value'],v['e2e']['frac_of_device_value'])
print(d.get('cpu_baseline'))
P
timeout 150 python bench.py --workload n1 > $OUT/bench_n1_prefix.json 2> $OUT/bench_n1_prefix.err; cut -c1-200 $OUT/bench_n1_prefix.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2>&1; cut -c1-300 $OUT/bench_reference.json
