#!/usr/bin/env bash
# r05, 8 GPUs: the driver's N=8 bench command (C2 weak scaling headline + extras: C3 strong scaling over 8 KV-head shards,
# C5 4 -> 4 KV transfer over CUDA IPC, all-reduce vs NCCL and the reference kernels) and the multi-GPU tests.
set -uo pipefail
OUT=gpurun_out/r05n4
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551"
timeout 600 $TR bench.py --gpus 4 --steps 20 --warmup 5 > $OUT/bench_n4.json 2> $OUT/bench_n4.err
echo "bench n8 exit $?"; grep -v "^\[W\|^$\|\*\*\*\|OMP_NUM\|^\[rank.*Warning" $OUT/bench_n4.err | tail -5 | cut -c1-300
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/r05n4/bench_n4.json').read().strip().splitlines()[-1])
    print('value',d['value'],'e2e',d['e2e']['value'],'errors',d.get('extra_errors'),'wall',d.get('wall_s'))
    x=d.get('extra',{})
    cs=x.get('c3_strong',{}); print('c3_strong',cs.get('value'),cs.get('efficiency_vs_one_gpu'),cs.get('e2e',{}).get('value'), cs.get('ms_per_step'))
    kv=x.get('kv_transfer',{}); print('kv bit_exact',kv.get('bit_exact'))
    for k,v in kv.get('cases',{}).items(): print('  ',k,v['value'],v.get('aggregate'),v.get('time_to_signal_ms'),v['bit_exact'],v.get('max_ctas'))
    ar=x.get('allreduce',{}); print('allreduce bit_exact',ar.get('bit_exact'))
    for k,v in ar.get('cases',{}).items(): print('  ',k,v['oneshot_us'],v['oneshot_registered_input_us'],v['nccl_us'],v.get('reference_cuda_us'),v.get('bit_identical_to_reference_cuda'))
except Exception as e: print('parse failed',e)
P
timeout 300 python -m pytest tests/test_kv_transfer_gpu.py tests/test_all_reduce_gpu.py tests/test_cache_gpu.py -q -k "two_gpu or two_process or other_gpu or multi_gpu or world" > $OUT/pytest_multi.log 2>&1
echo "pytest multi exit $?"; tail -3 $OUT/pytest_multi.log
