#!/usr/bin/env bash
# r05, 2 GPUs: KV-transfer put variants (SIMT warp chunks / cp.async.bulk ring) x CTA budget x layers per launch.
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_r05_kv2.sh'
set -uo pipefail
OUT=gpurun_out/r05
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
# loopback + peer parity of both variants first (a hang here must not take the sweep with it)
for v in 0 1; do
  B200_KV_PUT_VARIANT=$v timeout 300 python -m pytest tests/test_kv_transfer_gpu.py -x -q > $OUT/pytest_kv_variant$v.log 2>&1
  echo "pytest kv variant $v exit $?"; tail -3 $OUT/pytest_kv_variant$v.log
done
V1=""
grep -q "passed" $OUT/pytest_kv_variant1.log && ! grep -q "failed\|error" $OUT/pytest_kv_variant1.log && V1=1
SW=""
for c in 4 8 16 32 64 148 296; do
  for l in 1 4; do
    SW="$SW,0:$c:$l"
    [ -n "$V1" ] && SW="$SW,1:$c:$l"
  done
done
SW=${SW#,}
timeout 420 $TR tools/kv_transfer_bench.py --prompts 8 --iters 4 --modes bulk,per_layer,ref_api --sweep "$SW" \
    > $OUT/kv_put_sweep_p8.jsonl 2> $OUT/kv_put_sweep_p8.err
echo "sweep p8 exit $?"; tail -3 $OUT/kv_put_sweep_p8.err; wc -l $OUT/kv_put_sweep_p8.jsonl
SW1="0:16:1,0:32:1,0:32:4,0:64:4"
[ -n "$V1" ] && SW1="$SW1,1:8:1,1:16:1,1:16:4,1:32:4"
timeout 300 $TR tools/kv_transfer_bench.py --prompts 1 --iters 6 --modes bulk,per_layer,ref_api --sweep "$SW1" \
    > $OUT/kv_put_sweep_p1.jsonl 2> $OUT/kv_put_sweep_p1.err
echo "sweep p1 exit $?"; tail -3 $OUT/kv_put_sweep_p1.err
timeout 300 $TR tools/kv_transfer_bench.py --prompts 8 --iters 3 --modes paced,nccl --sweep "$SW1" \
    > $OUT/kv_paced_nccl_p8.jsonl 2> $OUT/kv_paced_nccl_p8.err
echo "paced/nccl exit $?"; tail -3 $OUT/kv_paced_nccl_p8.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r05/kv_*.jsonl')):
    print('==',f)
    for l in open(f):
        try: d=json.loads(l)
        except Exception: continue
        print(d['pattern'][:4],d['mode'],d.get('put_variant'),d.get('max_ctas'),d.get('layers_per_launch'),d['prompts_in_flight'],d['value'],d['best_GBps'],d.get('time_to_signal_ms'),d['bit_exact'])
P
