#!/usr/bin/env bash
# r05, 2 GPUs: the driver's N=2 bench command (extras: c3_strong, kv_transfer over CUDA IPC, allreduce vs the reference
# kernels), the chained per-layer put measurements, and the multi-GPU tests the 1-GPU suite skips.
set -uo pipefail
OUT=gpurun_out/r05g
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521"
timeout 300 python -m pytest tests/test_kv_transfer_gpu.py tests/test_all_reduce_gpu.py tests/test_cache_gpu.py -q -k "two_gpu or two_process or other_gpu or multi_gpu or world" > $OUT/pytest_multi.log 2>&1
echo "pytest multi exit $?"; tail -3 $OUT/pytest_multi.log
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $OUT/bench_n2.json 2> $OUT/bench_n2.err
echo "bench n2 exit $?"; tail -3 $OUT/bench_n2.err | cut -c1-300
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/r05g/bench_n2.json').read().strip().splitlines()[-1])
    print('value',d['value'],'e2e',d['e2e']['value'],'errors',d.get('extra_errors'))
    x=d.get('extra',{})
    cs=x.get('c3_strong',{}); print('c3_strong',cs.get('value'),cs.get('efficiency_vs_one_gpu'),cs.get('e2e',{}).get('value'))
    kv=x.get('kv_transfer',{}); print('kv bit_exact',kv.get('bit_exact'))
    for k,v in kv.get('cases',{}).items(): print('  ',k,v['value'],v.get('time_to_signal_ms'),v['bit_exact'],v.get('max_ctas'),v.get('put_variant'))
    ar=x.get('allreduce',{}); print('allreduce bit_exact',ar.get('bit_exact'))
    for k,v in ar.get('cases',{}).items(): print('  ',k,v['oneshot_us'],v['nccl_us'],v.get('reference_cuda_us'),v.get('bit_identical_to_reference_cuda'))
except Exception as e: print('parse failed',e)
P
timeout 300 $TR tools/kv_transfer_bench.py --prompts 8 --iters 4 --modes per_layer,ref_api --sweep "0:32:1,0:32:4,0:32:8,0:16:4" \
    > $OUT/kv_chained_p8.jsonl 2> $OUT/kv_chained_p8.err
timeout 300 $TR tools/kv_transfer_bench.py --prompts 8 --iters 4 --modes per_layer,ref_api --sweep "0:32:1,0:32:4" --comm-stream 0 \
    > $OUT/kv_unchained_p8.jsonl 2> $OUT/kv_unchained_p8.err
timeout 300 $TR tools/kv_transfer_bench.py --prompts 1 --iters 6 --modes per_layer,ref_api,bulk --sweep "0:32:1,0:32:4" \
    > $OUT/kv_chained_p1.jsonl 2> $OUT/kv_chained_p1.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r05g/kv_*.jsonl')):
    print('==',f)
    for l in open(f):
        try: d=json.loads(l)
        except Exception: continue
        print(d['pattern'][:4],d['mode'],d.get('put_variant'),d.get('max_ctas'),'lpl',d.get('layers_per_launch'),'side',d.get('comm_stream'),d['prompts_in_flight'],d['value'],d['best_GBps'],d['bit_exact'])
P
