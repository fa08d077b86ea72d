#!/usr/bin/env bash
# 2-GPU check of the final tree: the multi-GPU tests and the driver's bench command at N=2
set -uo pipefail
OUT=gpurun_out/r05n2f
mkdir -p $OUT
timeout 300 python -m pytest tests/test_kv_transfer_gpu.py tests/test_all_reduce_gpu.py -m gpu -q > $OUT/pytest_n2.log 2>&1
echo "pytest exit $?"; tail -3 $OUT/pytest_n2.log | cut -c1-200
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 500 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $OUT/bench_n2.json 2> $OUT/bench_n2.err
echo "bench n2 exit $?"; grep -v "^\[W\|^$\|\*\*\*\|OMP_NUM" $OUT/bench_n2.err | tail -3 | cut -c1-300
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/r05n2f/bench_n2.json').read().strip().splitlines()[-1])
    print('value',d['value'],'e2e',d['e2e']['value'],'errors',d.get('extra_errors'))
    x=d.get('extra',{})
    cs=x.get('c3_strong',{}); print('c3_strong',cs.get('value'),cs.get('efficiency_vs_one_gpu'))
    kv=x.get('kv_transfer',{}); print('kv bit_exact',kv.get('bit_exact'))
    for k,v in kv.get('cases',{}).items(): print('  ',k,v['value'],v.get('time_to_signal_ms'),v['bit_exact'])
    ar=x.get('allreduce',{}); print('allreduce bit_exact',ar.get('bit_exact'))
    for k,v in ar.get('cases',{}).items(): print('  ',k,v['oneshot_us'],v['nccl_us'],v.get('reference_cuda_us'),v.get('bit_identical_to_reference_cuda'))
except Exception as e: print('parse failed',e)
P
